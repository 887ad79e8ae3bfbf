"""Reference-produced vectors for IntegratedActionModelEuler(DifferentialActionModelLQR) (tests/golden/
ref_golden_difflqr.npz, written by tests/golden/make_ref_golden_difflqr.py from the reference's legacy NumPy classes):
the oracle's two-step statement (oracle/models.hpp ActionModelEulerDiffLQR) and the product's host-side mapping onto
the LQR kernels (crocoddyl_b200/integrated.py, evaluated here through the ORACLE's ActionModelLQR) must reproduce
xnext, Fx, Fu exactly as the reference computes them, and cost / Lx / Lu / Lxx / Lxu / Luu up to the time-step factor
the C++ integrator applies and the legacy one does not (euler.hxx:68,117-121 vs integrated_action.py:27,52-56)."""
import os

import numpy as np
import pytest

import crocoddyl_b200 as cb
import oracle_lib as ol
from integrated_cases import oracle_params

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_golden_difflqr.npz"))


def _case(name):
    pre = "euler_difflqr_%s/" % name
    return {k[len(pre):]: G[k] for k in G.files if k.startswith(pre)}


def _model(c):
    d = cb.DifferentialActionModelLQR(3, 2, bool(c["drift_free"]))
    for k in ("Fq", "Fv", "Fu", "f0", "Lxx", "Lxu", "Luu", "lx", "lu"):
        setattr(d, k, c["p_" + k][None])  # one problem: leading batch dimension of 1
    return d


@pytest.mark.parametrize("name", ["driftfree", "drift"])
@pytest.mark.parametrize("route", ["oracle_two_step", "mapping_on_lqr"])
def test_node_quantities_equal_the_reference(name, route):
    c = _case(name)
    dt = float(c["dt"])
    d = _model(c)
    assert float(c["cost_scaled"]) == 0.
    for x, u, i in zip(c["x"], c["u"], range(len(c["x"]))):
        if route == "oracle_two_step":
            ob = ol.OracleBatch("euler_difflqr", "ddp", 1, 6, 2, oracle_params(d, dt), x[None])
        else:
            ob = ol.OracleBatch("lqr", "ddp", 1, 6, 2, cb.IntegratedActionModelEuler(d, dt).params(), x[None])
        ob.set_candidate(np.stack([x, x])[None], u[None, None], False)
        ob.calc()
        ob.calcdiff()
        tol = dict(rtol=1e-12, atol=1e-13)
        np.testing.assert_allclose(ob.get("xnext", 0, 0), c["xnext"][i], **tol)
        np.testing.assert_allclose(ob.get("Fx", 0, 0), c["Fx"][i], **tol)
        np.testing.assert_allclose(ob.get("Fu", 0, 0), c["Fu"][i], **tol)
        np.testing.assert_allclose(ob.get("node_cost", 0, 0)[0], dt * c["cost"][i], **tol)
        for f in ("Lx", "Lu", "Lxx", "Lxu", "Luu"):
            np.testing.assert_allclose(ob.get(f, 0, 0), dt * c[f][i], err_msg=f, **tol)
