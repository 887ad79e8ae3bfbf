"""CPU tests of the oracle itself (the checker), modelled on the reference's own tests:
unittest/test_solvers.cpp:65-110 (DDP/FDDP vs SolverKKT, xs/us to 1e-9) and
unittest/test_actions.cpp:68-103 (calcDiff vs NumDiff, tol 3e4*sqrt(2 eps))."""
import numpy as np
import pytest

import oracle_lib as ol


def _lqr_random(n, m, rng, drift_free=True):
    G = rng.standard_normal((n, n))
    Fx = np.eye(n) + 0.1 * G / np.sqrt(n)
    Fu = rng.standard_normal((n, m))
    A = rng.standard_normal((n, n))
    Bm = rng.standard_normal((m, m))
    return ol.lqr_params(n, m, drift_free, Fx=Fx, Fu=Fu, f0=rng.standard_normal(n), Lxx=A.T @ A / n + np.eye(n),
                         Lxu=0.1 * rng.standard_normal((n, m)), Luu=Bm.T @ Bm / m + np.eye(m),
                         lx=rng.standard_normal(n), lu=rng.standard_normal(m))


@pytest.mark.parametrize("solver", ["ddp", "fddp"])
@pytest.mark.parametrize("case", ["lqr", "lqr_drift", "lqr_random", "unicycle"])
def test_solver_against_kkt(solver, case):
    T = 10
    rng = np.random.default_rng(3)
    if case == "unicycle":
        n, m, kind, params = 3, 2, "unicycle", ol.unicycle_params()
        x0 = np.array([[0.2, -0.1, 0.05]])  # reference test uses x0 = zero state (trivial); stay in the Newton basin
    else:
        n, m, kind = 20, 10, "lqr"
        if case == "lqr":
            params = ol.lqr_params(n, m, True)
        elif case == "lqr_drift":
            params = ol.lqr_params(n, m, False)
        else:
            params = _lqr_random(n, m, rng, False)
        x0 = np.zeros((1, n)) if case != "lqr_random" else rng.standard_normal((1, n))
    ob = ol.OracleBatch(kind, solver, T, n, m, params, x0)
    xs = np.repeat(x0[:, None, :], T + 1, 1)
    us = np.zeros((1, T, m))
    ob.solve(xs, us, 100, False, 1e-9)
    ok, kx, ku = ob.kkt_solve(0, xs[0], us[0], 100)
    assert ok == 1
    # LQR: one Newton step is exact -> the reference's 1e-9 (test_solvers.cpp:104-109). Unicycle from a
    # non-zero x0 (the reference's factory uses the trivial zero state): both stop at th_stop = 1e-9 on
    # ||Qu||^2, i.e. agree to ~sqrt(th_stop).
    tol = 1e-9 if kind == "lqr" else 1e-4
    np.testing.assert_allclose(ob.get_all("xs", 0), kx, atol=tol)
    np.testing.assert_allclose(ob.get_all("us", 0), ku, atol=tol)


def test_lqr_ddp_one_iteration_solves():
    """DDP solves an LQR problem in one Newton step (reference test idea)."""
    n, m, T = 12, 4, 20
    rng = np.random.default_rng(5)
    params = _lqr_random(n, m, rng)
    x0 = rng.standard_normal((1, n))
    ob = ol.OracleBatch("lqr", "ddp", T, n, m, params, x0)
    xs = np.repeat(x0[:, None, :], T + 1, 1)
    ob.solve(xs, np.zeros((1, T, m)), 100, False, 1e-9)
    tr = ob.trace(0)
    assert tr.shape[0] <= 3
    assert tr[0, 11] == 0  # alpha = 1 accepted at the first iteration


@pytest.mark.parametrize("kind", ["unicycle", "lqr"])
def test_calcdiff_against_numdiff(kind):
    rng = np.random.default_rng(11)
    if kind == "unicycle":
        n, m, params = 3, 2, ol.unicycle_params()
    else:
        n, m = 7, 3
        params = _lqr_random(n, m, rng, False)
    ob = ol.OracleBatch(kind, "ddp", 2, n, m, params, np.zeros((1, n)))
    tol = 3e4 * np.sqrt(2 * np.finfo(float).eps)  # unittest_common.hpp:20
    for _ in range(5):
        x, u = rng.standard_normal(n), rng.standard_normal(m)
        a, d = ob.numdiff(0, x, u)
        for k in ("Fx", "Fu", "Lx", "Lu"):
            np.testing.assert_allclose(a[k], d[k], atol=tol, rtol=1e-5)


def test_fddp_trace_fields():
    T, B = 30, 4
    rng = np.random.default_rng(0)
    x0 = np.stack([rng.uniform(-2, 2, B), rng.uniform(-2, 2, B), rng.uniform(-np.pi, np.pi, B)], 1)
    ob = ol.OracleBatch("unicycle", "fddp", T, 3, 2, ol.unicycle_params(), x0)
    xs = np.repeat(x0[:, None, :], T + 1, 1)
    n = ob.solve(xs, np.zeros((B, T, 2)), 100, False, 1e-9)
    assert n >= 1
    for b in range(B):
        tr = ob.trace(b)
        assert np.all(np.diff(tr[:, 0]) == 1)
        assert np.all(tr[:, 1] > 0)
