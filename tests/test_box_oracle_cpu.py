"""Pins the oracle's BoxQP / SolverBoxDDP / SolverBoxFDDP (oracle/box.hpp, restating src/core/solvers/box-qp.cpp,
box-ddp.cpp, box-fddp.cpp) with the reference's own test structures (unittest/test_boxqp.cpp:24-133) and with a
third party: scipy.optimize bounded solvers give the optimum of the same box QP."""
import numpy as np
import pytest
from scipy.optimize import lsq_linear, minimize

import oracle_lib as ol

INF = np.inf


@pytest.mark.parametrize("nx", [2, 3, 5])
def test_unconstrained_qp_with_identity_hessian(nx):
    rng = np.random.default_rng(nx)
    g, xinit = rng.uniform(-1, 1, nx), rng.uniform(-1, 1, nx)
    lb, ub = -INF * np.ones(nx), INF * np.ones(nx)
    sol = ol.boxqp_solve(np.eye(nx), g, lb, ub, xinit, reg=0.)
    np.testing.assert_allclose(sol["x"], -g, atol=1e-9)
    assert not sol["clamped"].any()
    reg = 10 ** rng.uniform(-9, 2)
    sol = ol.boxqp_solve(np.eye(nx), g, lb, ub, xinit, reg=reg)
    np.testing.assert_allclose(sol["x"], -g / (1 + reg), atol=1e-9)
    assert not sol["clamped"].any()


@pytest.mark.parametrize("nx", [2, 4, 5])
def test_unconstrained_qp(nx):
    rng = np.random.default_rng(10 + nx)
    A = rng.uniform(-1, 1, (nx, nx))
    H = A.T @ A
    H = 0.5 * (H + H.T) + 1e-3 * np.eye(nx)
    g, xinit = rng.uniform(-1, 1, nx), rng.uniform(-1, 1, nx)
    lb, ub = -INF * np.ones(nx), INF * np.ones(nx)
    sol = ol.boxqp_solve(H, g, lb, ub, xinit, reg=0.)
    np.testing.assert_allclose(sol["x"], -np.linalg.solve(H, g), rtol=1e-8, atol=1e-9)
    np.testing.assert_allclose(sol["Hff_inv"], np.linalg.inv(H), rtol=1e-7, atol=1e-9)
    reg = 10 ** rng.uniform(-9, 2)
    sol = ol.boxqp_solve(H, g, lb, ub, xinit, reg=reg)
    np.testing.assert_allclose(sol["x"], -np.linalg.solve(H + reg * np.eye(nx), g), rtol=1e-8, atol=1e-9)
    assert not sol["clamped"].any()


@pytest.mark.parametrize("nx", [2, 3, 5])
def test_box_qp_with_identity_hessian(nx):
    rng = np.random.default_rng(20 + nx)
    g = rng.uniform(-1, 1, nx)
    lb, ub = np.zeros(nx), np.ones(nx)
    sol = ol.boxqp_solve(np.eye(nx), g, lb, ub, rng.uniform(-1, 1, nx), reg=0.)
    want = np.clip(-g, lb, ub)
    np.testing.assert_allclose(sol["x"], want, atol=1e-9)
    assert np.array_equal(sol["clamped"], want != -g)
    assert sol["Hff_inv"].shape == ((want == -g).sum(),) * 2


def test_negative_definite_hessian_is_a_backward_error():
    with pytest.raises(ArithmeticError):
        ol.boxqp_solve(-np.eye(3), np.ones(3), -np.ones(3), np.ones(3), np.zeros(3), reg=0.)


@pytest.mark.parametrize("seed", range(8))
def test_box_qp_optimum_matches_scipy(seed):
    """third-party pin: min 1/2 x'Hx + q'x, lb <= x <= ub  ==  bounded least squares with H = R'R, q = -R'b."""
    rng = np.random.default_rng(100 + seed)
    n = int(rng.integers(2, 8))
    R = rng.standard_normal((n + 2, n))
    H = R.T @ R
    b = 2.0 * rng.standard_normal(n + 2)
    q = -R.T @ b
    lb = -rng.uniform(0.05, 1.0, n)
    ub = rng.uniform(0.05, 1.0, n)
    sol = ol.boxqp_solve(H, q, lb, ub, np.zeros(n), th_grad=1e-12, reg=0.)
    ref = lsq_linear(R, b, bounds=(lb, ub), method="bvls", tol=1e-14)
    np.testing.assert_allclose(sol["x"], ref.x, atol=1e-7)
    f = lambda x: 0.5 * x @ H @ x + q @ x
    assert f(sol["x"]) <= f(ref.x) + 1e-10
    # clamped set = variables on a bound with the gradient pushing outwards; Hff_inv = inverse of the free block
    free = ~sol["clamped"]
    np.testing.assert_allclose(sol["Hff_inv"], np.linalg.inv(H[np.ix_(free, free)]), rtol=1e-7, atol=1e-9)
    assert np.all((sol["x"][sol["clamped"]] == lb[sol["clamped"]]) | (sol["x"][sol["clamped"]] == ub[sol["clamped"]]))


def _lqr_case(seed, n=4, m=3, T=12, B=3):
    rng = np.random.default_rng(seed)
    P = []
    for b in range(B):
        A, Bm = rng.standard_normal((n, n)), rng.standard_normal((m, m))
        P.append(ol.lqr_params(n, m, False, Fx=np.eye(n) + 0.1 * rng.standard_normal((n, n)), Fu=rng.standard_normal((n, m)),
                               f0=0.1 * rng.standard_normal(n), Lxx=A.T @ A / n + np.eye(n), Lxu=0.1 * rng.standard_normal((n, m)),
                               Luu=Bm.T @ Bm / m + 0.1 * np.eye(m), lx=rng.standard_normal(n), lu=rng.standard_normal(m)))
    return np.stack(P), rng.standard_normal((B, n)), n, m, T, B


@pytest.mark.parametrize("kind", ["boxddp", "boxfddp"])
def test_box_solvers_without_limits_equal_the_plain_solvers(kind):
    P, x0, n, m, T, B = _lqr_case(1)
    xs, us = np.repeat(x0[:, None, :], T + 1, 1), np.zeros((B, T, m))
    a = ol.OracleBatch("lqr", kind, T, n, m, P, x0)
    b = ol.OracleBatch("lqr", kind[3:], T, n, m, P, x0)
    b.set("th_stop", 5e-5)  # box-ddp.cpp:29
    a.solve(xs, us, 20, False, 1e-9)
    b.solve(xs, us, 20, False, 1e-9)
    assert np.array_equal(a.scalars()["n_records"], b.scalars()["n_records"])
    np.testing.assert_array_equal(a.get_all("xs"), b.get_all("xs"))
    np.testing.assert_array_equal(a.get_all("us"), b.get_all("us"))


@pytest.mark.parametrize("kind", ["boxddp", "boxfddp"])
def test_box_solvers_respect_limits_and_reach_the_bounded_optimum(kind):
    """LQR with control limits is a convex QP in us: the control-limited solvers must end inside the box at a cost no
    worse than scipy's L-BFGS-B on the same problem (condensed through the rollout) and never above the clamped
    unconstrained solution."""
    P, x0, n, m, T, B = _lqr_case(2)
    lim = 0.15
    lb, ub = -lim * np.ones((B, m)), lim * np.ones((B, m))
    xs, us = np.repeat(x0[:, None, :], T + 1, 1), np.zeros((B, T, m))
    s = ol.OracleBatch("lqr", kind, T, n, m, P, x0)
    s.set_u_bounds(lb, ub)
    s.solve(xs, us, 100, False, 1e-9)
    sc = s.scalars()
    U = s.get_all("us")
    assert np.all(U >= -lim - 1e-12) and np.all(U <= lim + 1e-12)
    assert (np.abs(np.abs(U) - lim) < 1e-12).any(), "no active bound: the case does not exercise the box QP"
    plain = ol.OracleBatch("lqr", "fddp", T, n, m, P, x0)
    for b in range(B):
        def total_cost(uflat, b=b):
            one = ol.OracleBatch("lqr", "ddp", T, n, m, P[b], x0[b:b + 1])
            u = uflat.reshape(1, T, m)
            one.set_candidate(xs[b:b + 1], u, False)
            one.calc()
            X = np.concatenate([x0[b:b + 1], np.stack([one.get("xnext", 0, t) for t in range(T)])])
            one.set_candidate(X[None], u, True)
            return one.calc()[0]
        ref = minimize(total_cost, np.zeros(T * m), method="L-BFGS-B", bounds=[(-lim, lim)] * (T * m),
                       options=dict(maxiter=500, ftol=1e-15, gtol=1e-10))
        assert sc["cost"][b] <= ref.fun * (1 + 1e-6) + 1e-9, (sc["cost"][b], ref.fun)
        # feasible, and the reported cost is the cost of the rollout of us
        assert abs(total_cost(U[b].ravel()) - sc["cost"][b]) <= 1e-9 * max(1.0, abs(sc["cost"][b]))
