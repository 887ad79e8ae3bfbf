"""Regenerates tests/golden/*.npz from the CPU oracle:  python tests/golden/make_golden.py
(needs only the host: crocoddyl_b200's model classes are plain Python, the oracle is oracle/_build/liboracle.so)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_lib as ol  # noqa: E402
import crocoddyl_b200 as cb  # noqa: E402
from golden import cases  # noqa: E402


def oracle_run(name):
    cs = cases.build(name, cb)
    mod, x0, T = cs["model"], cs["x0"], cs["T"]
    params = mod.params()
    us = cs["us"]
    if us is None:  # quasiStatic(x0) as benchmark/arm-manipulation-optctrl.cpp:41-47
        u0 = np.stack([ol.arm_op(params, "quasistatic", x0[b, :mod.nu], mod.nu) for b in range(x0.shape[0])])
        us = np.repeat(u0[:, None, :], T, 1)
    ob = ol.OracleBatch(cs["kind"], cs["solver"], T, mod.nx, mod.nu, params, x0)
    ob.solve(cs["xs"], us, cs["maxiter"], False, cs["reginit"])
    sc = ob.scalars()
    out = {"init_us": us}
    for k in ("cost", "stop", "iter", "converged", "n_records", "xreg", "steplength", "is_feasible"):
        out["s_" + k] = sc[k]
    for f in cases.FIELDS:
        out[f] = ob.get_all(f)
    for b in range(x0.shape[0]):
        out["trace_%d" % b] = ob.trace(b)
    return out


def main():
    for name in cases.CASES:
        out = oracle_run(name)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, "iters", out["s_iter"], "cost", out["s_cost"])


if __name__ == "__main__":
    main()
