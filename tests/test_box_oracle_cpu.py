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


def _lqr_case(seed, n=4, m=2, T=15, B=3, fu=0.3):
    """-> params [B, np], x0 [B, n], n, m, T, B, list of the model matrices (for the independent NumPy cost below)."""
    rng = np.random.default_rng(seed)
    P, M = [], []
    for b in range(B):
        A, Bm = rng.standard_normal((n, n)), rng.standard_normal((m, m))
        d = dict(Fx=np.eye(n) + 0.05 * rng.standard_normal((n, n)), Fu=fu * rng.standard_normal((n, m)),
                 f0=0.02 * rng.standard_normal(n), Lxx=A.T @ A / n + np.eye(n), Lxu=0.05 * rng.standard_normal((n, m)),
                 Luu=Bm.T @ Bm / m + 0.5 * np.eye(m), lx=0.1 * rng.standard_normal(n), lu=0.1 * rng.standard_normal(m))
        M.append(d)
        P.append(ol.lqr_params(n, m, False, **d))
    return np.stack(P), rng.standard_normal((B, n)), n, m, T, B, M


def _cost_and_gradient(uflat, d, x0, T, m):
    """Total cost of the rollout of us on an LQR problem and its gradient by the adjoint recursion — plain NumPy,
    independent of the oracle (the objective handed to scipy)."""
    U = uflat.reshape(T, m)
    x, c, xs = x0, 0., [x0]
    for t in range(T):
        u = U[t]
        c += 0.5 * x @ d["Lxx"] @ x + 0.5 * u @ d["Luu"] @ u + x @ d["Lxu"] @ u + d["lx"] @ x + d["lu"] @ u
        x = d["Fx"] @ x + d["Fu"] @ u + d["f0"]
        xs.append(x)
    c += 0.5 * x @ d["Lxx"] @ x + d["lx"] @ x
    lam, g = d["Lxx"] @ x + d["lx"], np.zeros((T, m))
    for t in range(T - 1, -1, -1):
        x, u = xs[t], U[t]
        g[t] = d["Luu"] @ u + d["Lxu"].T @ x + d["lu"] + d["Fu"].T @ lam
        lam = d["Lxx"] @ x + d["Lxu"] @ u + d["lx"] + d["Fx"].T @ lam
    return c, g.ravel()


@pytest.mark.parametrize("kind", ["boxddp", "boxfddp"])
def test_box_solvers_without_limits_equal_the_plain_solvers(kind):
    P, x0, n, m, T, B, _ = _lqr_case(1)
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
@pytest.mark.parametrize("fu,lim", [(0.3, 0.3), (0.3, 0.6), (1.0, 0.15)])
def test_box_solvers_reach_the_bounded_optimum_of_scipy(kind, fu, lim):
    """Third-party pin at the solver level: an LQR problem with control limits is a convex box QP in us. The
    control-limited solvers must end inside the box, with active bounds, at the optimum scipy's L-BFGS-B finds for the
    same problem (objective + adjoint gradient written independently in NumPy above)."""
    P, x0, n, m, T, B, M = _lqr_case(3, fu=fu)
    lb, ub = -lim * np.ones((B, m)), lim * np.ones((B, m))
    xs, us = np.repeat(x0[:, None, :], T + 1, 1), np.zeros((B, T, m))
    s = ol.OracleBatch("lqr", kind, T, n, m, P, x0)
    s.set_u_bounds(lb, ub)
    s.solve(xs, us, 100, False, 1e-9)
    sc = s.scalars()
    assert np.all(sc["converged"] == 1)
    U = s.get_all("us")
    assert np.all(U >= -lim) and np.all(U <= lim)
    assert (np.abs(U) == lim).mean() > 0.1, "too few active bounds: the case does not exercise the box QP"
    for b in range(B):
        ref = minimize(_cost_and_gradient, np.zeros(T * m), args=(M[b], x0[b], T, m), jac=True, method="L-BFGS-B",
                       bounds=[(-lim, lim)] * (T * m), options=dict(maxiter=5000, ftol=1e-16, gtol=1e-12))
        mine = _cost_and_gradient(U[b].ravel(), M[b], x0[b], T, m)[0]
        np.testing.assert_allclose(sc["cost"][b], mine, rtol=1e-12)        # reported cost = cost of the rollout of us
        np.testing.assert_allclose(mine, ref.fun, rtol=1e-9)              # = the bounded optimum
        np.testing.assert_allclose(U[b].ravel(), ref.x, atol=2e-5)
