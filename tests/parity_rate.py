"""Test-infrastructure script (not collected by pytest): how often does the CUDA unicycle FDDP path take exactly the
reference's iteration count / accept sequence on config C2 (all 65,536 problems by default)? Compares with the CPU
oracle (pinned to the reference by tests/test_ref_golden_cpu.py) and writes a JSON record (committed under profiles/).

  python tests/parity_rate.py [N] [out.json]          # default library (FMA contraction, reciprocal pivots)
  CROCODDYL_B200_LIB=crocoddyl_b200/lib/libcrocoddyl_b200_parity.so python tests/parity_rate.py ...   # parity build
"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import crocoddyl_b200 as cb  # noqa: E402
import oracle_lib as ol  # noqa: E402


def run(N=65536, T=100, nsample=2048):
    rng = np.random.default_rng(0)
    x0 = np.stack([rng.uniform(-2, 2, 65536), rng.uniform(-2, 2, 65536), rng.uniform(-np.pi, np.pi, 65536)], 1)[:N]
    xs = np.repeat(x0[:, None, :], T + 1, 1)
    us = np.zeros((N, T, 2))
    m = cb.ActionModelUnicycle()
    p = cb.ShootingProblem(x0, [m] * T, m)
    s = cb.SolverFDDP(p)
    s.setStorage(False)
    s.enableTrace(100)
    conv = s.solve(xs, us, 100, False, 1e-9)
    sg = s.scalars()
    ob = ol.OracleBatch("unicycle", "fddp", T, 3, 2, ol.unicycle_params(), x0, fast=False)
    ob.solve(xs, us, 100, False, 1e-9)
    so = ob.scalars()
    same = so["n_records"] == sg["n_records"]
    # exact accept sequence (accepted alpha index, xreg, feasibility per iteration) where the counts agree
    seq_same = same.copy()
    first_div = {}
    check = np.where(same)[0] if N <= 8192 else np.concatenate([np.where(~same)[0], np.where(same)[0][:8192]])
    for b in np.where(~same)[0]:
        tg, to = s.trace(int(b)), ob.trace(int(b))
        n = min(len(tg), len(to))
        neq = np.where(np.any(tg[:n][:, (5, 7, 8, 11)] != to[:n][:, (5, 7, 8, 11)], axis=1))[0]
        first_div[int(b)] = int(neq[0]) if len(neq) else n
    nseq = 0
    for b in np.where(same)[0][:8192]:
        tg, to = s.trace(int(b)), ob.trace(int(b))
        ok = np.array_equal(tg[:, (5, 7, 8, 11)], to[:, (5, 7, 8, 11)])
        seq_same[b] = ok
        nseq += 1
    idx = np.random.default_rng(1).choice(N, min(nsample, N), replace=False)
    xg = s.get("xs")[idx]
    xo = np.stack([ob.get_all("xs", int(b)) for b in idx])
    err = np.abs(xg - xo).reshape(len(idx), -1).max(1) / np.maximum(1., np.abs(xo).reshape(len(idx), -1).max(1))
    oks = same[idx] & (so["converged"][idx] == 1)
    rec = {
        "lib": os.path.basename(cb._lib.LIB_PATH), "config": "C2 unicycle FDDP T=100 seed 0", "N": int(N),
        "same_iteration_count_frac": float(same.mean()), "n_differ": int((~same).sum()),
        "same_frac_among_oracle_converged_lt100": float(same[(so["converged"] == 1) & (so["n_records"] < 100)].mean()),
        "accept_sequence_checked": int(nseq), "accept_sequence_identical_frac": float(seq_same[np.where(same)[0][:8192]].mean()),
        "converged_frac_gpu": float(np.mean(conv)), "converged_frac_oracle": float(so["converged"].mean()),
        "converged_flags_equal_frac": float((so["converged"].astype(bool) == conv).mean()),
        "mean_iters_oracle": float(so["n_records"].mean()), "mean_iters_gpu": float(sg["n_records"].mean()),
        "cost_rel_err_max_where_same": float(np.max(np.abs(sg["cost"][same] - so["cost"][same]) / np.abs(so["cost"][same]))),
        "xs_sample": int(len(idx)), "xs_rel_err_max_where_same_converged": float(err[oks].max()),
        "xs_rel_err_le_1e-9_frac": float((err[oks] <= 1e-9).mean()), "xs_rel_err_le_1e-8_frac": float((err[oks] <= 1e-8).mean()),
        "differing_problems": [{"b": int(b), "oracle_iters": int(so["n_records"][b]), "gpu_iters": int(sg["n_records"][b]),
                                "first_diverging_iteration": first_div[int(b)]} for b in np.where(~same)[0][:64]],
        "kernel_ms": float(s.timing()["kernel_ms"]), "oracle_seconds": float(ob.seconds),
    }
    return rec


if __name__ == "__main__":
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
    rec = run(N)
    print(json.dumps(rec))
    if len(sys.argv) > 2:
        with open(sys.argv[2], "w") as fh:
            json.dump(rec, fh, indent=1)
