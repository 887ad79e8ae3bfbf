"""CPU-only experiment (test infrastructure): how many C2 unicycle FDDP solves change their iteration count when the
results of sin / cos move by ONE ulp on a quarter of the calls? Oracle (glibc libm) vs the same oracle built with
-DORACLE_PERTURB_SINCOS (oracle/models.hpp libm_result). This is the floor under which no implementation with a
different libm (CUDA's sin/cos on the device) can match the reference's iteration counts on every problem."""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as ol  # noqa: E402


def run(N, T=100):
    rng = np.random.default_rng(0)
    x0 = np.stack([rng.uniform(-2, 2, 65536), rng.uniform(-2, 2, 65536), rng.uniform(-np.pi, np.pi, 65536)], 1)[:N]
    xs = np.repeat(x0[:, None, :], T + 1, 1)
    us = np.zeros((N, T, 2))
    res = []
    for name in ("liboracle.so", "liboracle_perturb.so"):
        real_load = ol.load
        ol.load = lambda fast=False, _n=name: _load(_n)
        ob = ol.OracleBatch("unicycle", "fddp", T, 3, 2, ol.unicycle_params(), x0)
        ol.load = real_load
        ob.solve(xs, us, 100, False, 1e-9)
        res.append(ob.scalars())
    a, b = res
    same = a["n_records"] == b["n_records"]
    d = np.where(~same)[0]
    return {"experiment": "oracle vs oracle with sin/cos moved by 1 ulp on 1/4 of the calls", "N": int(N),
            "same_iteration_count_frac": float(same.mean()), "n_differ": int((~same).sum()),
            "iters_of_differing_min": int(a["n_records"][d].min()) if len(d) else None,
            "cost_rel_err_max_where_same": float(np.max(np.abs(a["cost"][same] - b["cost"][same]) / np.abs(a["cost"][same]))),
            "differing": [(int(i), int(a["n_records"][i]), int(b["n_records"][i])) for i in d[:32]]}


def _load(name):
    lib = C.CDLL(os.path.join(ROOT, "oracle", "_build", name))
    base = ol.__dict__["_base_load"](False)
    for k in dir(base):
        pass
    return _typed(lib)


def _typed(lib):
    # same prototypes as oracle_lib.load
    ref = ol._base_load(False)
    for name in ("orc_batch_create", "orc_batch_destroy", "orc_batch_set", "orc_batch_set_alphas", "orc_batch_solve",
                 "orc_batch_set_candidate", "orc_batch_calc", "orc_batch_calcdiff", "orc_batch_backward",
                 "orc_batch_try_step", "orc_batch_expected_improvement", "orc_get", "orc_batch_scalars", "orc_get_trace"):
        f, g = getattr(lib, name), getattr(ref, name)
        f.argtypes, f.restype = g.argtypes, g.restype
    return lib


if __name__ == "__main__":
    ol._base_load = ol.load
    rec = run(int(sys.argv[1]) if len(sys.argv) > 1 else 65536)
    print(json.dumps(rec))
    if len(sys.argv) > 2:
        json.dump(rec, open(sys.argv[2], "w"), indent=1)
