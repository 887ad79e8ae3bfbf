"""Golden vectors produced by the REFERENCE for SURVEY 8 f3's DifferentialActionModelLQR under the Euler integrator:
python tests/golden/make_ref_golden_difflqr.py  (build container only; reads /root/reference, writes
tests/golden/ref_golden_difflqr.npz; nothing in tests/ reads the reference at run time).

Executed unmodified: the legacy pure-NumPy crocoddyl of the reference tree,
  /root/reference/unittest/python/crocoddyl/differential_action.py:189-244 (DifferentialActionModelLQR) and
  /root/reference/unittest/python/crocoddyl/integrated_action.py:4-58 (IntegratedActionModelEuler),
with the import stubs of make_ref_golden.py. One known difference from the C++ (recorded in the file as `cost_scaled =
0`): the legacy integrator returns the differential cost and its derivatives UNSCALED (integrated_action.py:27,52-56),
the C++ multiplies them by the time step (euler.hxx:68,117-121). Dynamics (xnext, Fx, Fu) are the same statement."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_ref_golden as g  # noqa: E402  (stubs, numpy shims)


def main():
    sys.meta_path.insert(0, g._StubFinder())
    sys.path.insert(0, os.path.join(g.REF, "unittest", "python"))
    import crocoddyl as lc

    out = {}
    rng = np.random.default_rng(424242)
    for name, drift_free, dt in (("driftfree", True, 0.05), ("drift", False, 0.02)):
        np.random.seed(11 if drift_free else 12)  # the legacy constructor draws its matrices from numpy's global state
        dam = lc.DifferentialActionModelLQR(3, 2, driftFree=drift_free)
        iam = lc.IntegratedActionModelEuler(dam, timeStep=dt, withCostResiduals=False)  # the LQR model has no residuals (no ncost)
        d = iam.createData()
        X, U = rng.standard_normal((8, 6)), rng.standard_normal((8, 2))
        rec = {k: [] for k in ("cost", "xnext", "Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu")}
        for x, u in zip(X, U):
            iam.calcDiff(d, x, u)
            rec["cost"].append(float(d.cost))
            for k in rec:
                if k != "cost":
                    rec[k].append(np.array(getattr(d, k), dtype=np.float64))
        case = dict(x=X, u=U, dt=np.array(dt), drift_free=np.array(float(drift_free)), cost_scaled=np.array(0.))
        for k in ("Fq", "Fv", "Fu", "f0", "Lxx", "Lxu", "Luu", "lx", "lu"):
            case["p_" + k] = np.array(getattr(dam, k), dtype=np.float64)
        for k, v in rec.items():
            case[k] = np.stack([np.asarray(a, dtype=np.float64) for a in v])
        for k, v in case.items():
            out["euler_difflqr_%s/%s" % (name, k)] = v
    np.savez_compressed(os.path.join(HERE, "ref_golden_difflqr.npz"), **out)
    print("wrote", len(out), "arrays")


if __name__ == "__main__":
    main()
