"""Run in a subprocess by tests/test_parity_build_gpu.py with CROCODDYL_B200_LIB pointing at the PARITY build of the
library (python -m crocoddyl_b200.build --parity: -fmad=false, IEEE divisions by the LLT pivots, sqrt). Test
infrastructure. Prints one JSON object: for each case whether the CUDA results are BIT-identical to the CPU oracle's
(built -ffp-contract=off)."""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
import crocoddyl_b200 as cb  # noqa: E402
import oracle_lib as ol  # noqa: E402

FIELDS = ("xs", "us", "K", "k", "Vx", "Vxx")


def lqr_case(n, m, B, T, kind, maxiter=20):
    rng = np.random.default_rng(n)
    mod = cb.ActionModelLQR(n, m, False)
    mod.Fx = np.eye(n) + 0.1 * rng.standard_normal((B, n, n)) / np.sqrt(n)
    mod.Fu = rng.standard_normal((B, n, m))
    A, Bm = rng.standard_normal((B, n, n)), rng.standard_normal((B, m, m))
    mod.Lxx = np.einsum("bki,bkj->bij", A, A) / n + np.eye(n)
    mod.Luu = np.einsum("bki,bkj->bij", Bm, Bm) / m + np.eye(m)
    mod.Lxu = 0.1 * rng.standard_normal((B, n, m))
    mod.lx, mod.lu, mod.f0 = rng.standard_normal((B, n)), rng.standard_normal((B, m)), rng.standard_normal((B, n))
    x0 = rng.standard_normal((B, n))
    problem = cb.ShootingProblem(x0, [mod] * T, mod)
    solver = (cb.SolverFDDP if kind == "fddp" else cb.SolverDDP)(problem)
    ob = ol.OracleBatch("lqr", kind, T, n, m, mod.params(), x0)
    xs, us = np.repeat(x0[:, None, :], T + 1, 1), np.zeros((B, T, m))
    solver.solve(xs, us, maxiter, False, 1e-9)
    ob.solve(xs, us, maxiter, False, 1e-9)
    out = {f: bool(np.array_equal(solver.get(f), ob.get_all(f))) for f in FIELDS}
    out["max_abs_diff"] = float(max(np.max(np.abs(solver.get(f) - ob.get_all(f))) for f in FIELDS))
    sg, so = solver.scalars(), ob.scalars()
    out["decisions"] = bool(all(np.array_equal(sg[k], so[k]) for k in ("n_records", "iter", "xreg", "ureg", "steplength", "is_feasible")))
    out["cost_rel_diff"] = float(np.max(np.abs(sg["cost"] - so["cost"]) / np.abs(so["cost"])))
    return out


def unicycle_case(B, T):
    rng = np.random.default_rng(0)
    x0 = np.stack([rng.uniform(-2, 2, B), rng.uniform(-2, 2, B), rng.uniform(-np.pi, np.pi, B)], 1)
    m = cb.ActionModelUnicycle()
    solver = cb.SolverFDDP(cb.ShootingProblem(x0, [m] * T, m))
    xs, us = np.repeat(x0[:, None, :], T + 1, 1), np.zeros((B, T, 2))
    solver.solve(xs, us, 100, False, 1e-9)
    ob = ol.OracleBatch("unicycle", "fddp", T, 3, 2, ol.unicycle_params(), x0)
    ob.solve(xs, us, 100, False, 1e-9)
    g, o = solver.get("xs"), ob.get_all("xs")
    return {"bitwise_equal_problems": int(sum(np.array_equal(g[b], o[b]) for b in range(B))), "B": B,
            "same_iterations": int((solver.scalars()["n_records"] == ob.scalars()["n_records"]).sum()),
            "max_abs_diff": float(np.abs(g - o).max())}


if __name__ == "__main__":
    res = {"lib": os.path.basename(cb._lib.LIB_PATH)}
    for n, m in ((6, 2), (12, 4), (16, 8)):
        for kind in ("ddp", "fddp"):
            res["lqr_%d_%d_%s" % (n, m, kind)] = lqr_case(n, m, 4, 10, kind)
    res["lqr_36_12_fddp"] = lqr_case(36, 12, 2, 10, "fddp")
    res["unicycle"] = unicycle_case(256, 20)
    print(json.dumps(res))
