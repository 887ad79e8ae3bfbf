"""GPU parity tests of SolverBoxDDP / SolverBoxFDDP (BoxQP in computeGains, clamp in forwardPass; src/core/solvers/
box-ddp.cpp, box-fddp.cpp, box-qp.cpp) against the CPU oracle (oracle/box.hpp, pinned by tests/test_box_oracle_cpu.py),
through the C ABI, on all three device cores: thread-per-problem (unicycle), warp-per-problem (small LQR, arm) and
CTA-per-problem (LQR n >= 17)."""
import numpy as np
import pytest

import crocoddyl_b200 as cb
import oracle_lib as ol

pytestmark = pytest.mark.gpu
TOL = dict(rtol=1e-9, atol=1e-9)


def _lqr(B, n, m, seed, fu=0.3):
    rng = np.random.default_rng(seed)
    mod = cb.ActionModelLQR(n, m, False)
    mod.Fx = np.eye(n) + 0.05 * rng.standard_normal((B, n, n))
    mod.Fu = fu * rng.standard_normal((B, n, m))
    mod.f0 = 0.02 * rng.standard_normal((B, n))
    A, Bm = rng.standard_normal((B, n, n)), rng.standard_normal((B, m, m))
    mod.Lxx = np.einsum("bki,bkj->bij", A, A) / n + np.eye(n)
    mod.Luu = np.einsum("bki,bkj->bij", Bm, Bm) / m + 0.5 * np.eye(m)
    mod.Lxu = 0.05 * rng.standard_normal((B, n, m))
    mod.lx, mod.lu = 0.1 * rng.standard_normal((B, n)), 0.1 * rng.standard_normal((B, m))
    return mod, rng.standard_normal((B, n))


def _uni(B, seed):
    rng = np.random.default_rng(seed)
    x0 = np.stack([rng.uniform(-2, 2, B), rng.uniform(-2, 2, B), rng.uniform(-np.pi, np.pi, B)], 1)
    return cb.ActionModelUnicycle(), x0


def _arm(B, seed, nv=7):
    # a horizon long enough (20 x 10 ms) and a goal cost strong enough for the torques to reach their limits
    mod = cb.ArmActionModel(cb.synthetic_arm(nv), 1e-2, w_gripperPose=10., w_xReg=1e-3, w_uReg=1e-4)
    rng = np.random.default_rng(seed)
    return mod, np.concatenate([rng.uniform(-1, 1, (B, nv)), np.zeros((B, nv))], 1)


CASES = {
    # name: (family, builder, T, bounds, reginit, maxiter)
    "unicycle": ("unicycle", lambda: _uni(48, 1), 30, 0.8, 1e-9, 100),
    "lqr_small": ("lqr", lambda: _lqr(9, 6, 3, 2), 15, 0.3, 1e-9, 50),
    "lqr_large": ("lqr", lambda: _lqr(5, 20, 6, 3), 10, 0.3, 1e-9, 50),
    "arm": ("arm", lambda: _arm(6, 4), 20, None, 0.1, 60),
}


def _setup(name, kind):
    family, build, T, lim, reginit, maxiter = CASES[name]
    mod, x0 = build()
    B = x0.shape[0]
    term = mod.terminal_model if family == "arm" else mod
    problem = cb.ShootingProblem(x0, [mod] * T, term)
    if family == "arm":
        u0 = problem.quasiStatic(x0)
        us = np.repeat(u0[:, None, :], T, 1)
        lb, ub = u0.min(0) - 5.0, u0.max(0) + 5.0  # around the gravity torques: ~3/4 of the controls end on a limit
    else:
        us = np.zeros((B, T, mod.nu))
        lb, ub = -lim * np.ones(mod.nu), lim * np.ones(mod.nu)
    problem.set_u_bounds(lb, ub)
    solver = (cb.SolverBoxFDDP if kind == "boxfddp" else cb.SolverBoxDDP)(problem)
    ob = ol.OracleBatch(family, kind, T, mod.nx, mod.nu, mod.params(), x0)
    ob.set_u_bounds(lb, ub)
    assert solver.th_stop == 5e-5  # box-ddp.cpp:29
    if family == "arm":  # the default tolerance of the box solvers stops this case after two iterations
        solver.th_stop = 1e-7
        ob.set("th_stop", 1e-7)
    xs = np.repeat(x0[:, None, :], T + 1, 1)
    return problem, solver, ob, xs, us, lb, ub, reginit, maxiter


@pytest.mark.parametrize("kind", ["boxddp", "boxfddp"])
@pytest.mark.parametrize("name", ["unicycle", "lqr_small", "lqr_large", "arm"])
def test_stepwise_box_gains_and_clamped_rollout(name, kind):
    """computeDirection on a FEASIBLE candidate (the box QP runs only then, box-ddp.cpp:48) with controls on and near
    the limits, then tryStep: K, k, Quu_inv, the clamped Qu, Vx, Vxx, xs_try, us_try against the oracle."""
    problem, solver, ob, xs, us, lb, ub, reginit, maxiter = _setup(name, kind)
    rng = np.random.default_rng(5)
    us = np.clip(us + 0.6 * (ub - lb) * rng.standard_normal(us.shape), lb, ub)
    xs = problem.rollout(us)
    for s in (0, 1):  # the second pass warm-starts the QP from the first pass's k (box-ddp.cpp:57)
        if s == 0:
            solver.setCandidate(xs, us, True)
            ob.set_candidate(xs, us, True)
            solver.x_reg = solver.u_reg = 1e-6
            ob.set("xreg", 1e-6)
            ob.set("ureg", 1e-6)
        solver.computeDirection(True)
        ob.calcdiff()
        assert not ob.backward().any()
        for f in ("K", "k", "Qu", "Quu", "Vx", "Vxx"):
            np.testing.assert_allclose(solver.get(f), ob.get_all(f), err_msg="%s pass %d" % (f, s), **TOL)
        assert (solver.get("Qu") == 0.).any(), "no clamped control: the case does not exercise the box QP"
        Qi = solver.get("Quu_inv")
        np.testing.assert_allclose(Qi, np.transpose(Qi, (0, 1, 3, 2)), rtol=1e-9, atol=1e-12)  # inverse of the free block
    np.testing.assert_allclose(solver.stoppingCriteria(), ob.expected_improvement()[1], **TOL)
    for alpha in (1.0, 0.25):
        dV = solver.tryStep(alpha)
        dVo, st = ob.try_step(alpha)
        assert not st.any()
        np.testing.assert_allclose(dV, dVo, **TOL)
        ut = solver.get("us_try")
        np.testing.assert_allclose(ut, ob.get_all("us_try"), **TOL)
        np.testing.assert_allclose(solver.get("xs_try"), ob.get_all("xs_try"), **TOL)
        assert np.all(ut >= lb - 1e-15) and np.all(ut <= ub + 1e-15)
        assert ((ut == lb) | (ut == ub)).any()
        np.testing.assert_allclose(solver.expectedImprovement(), ob.expected_improvement()[0], **TOL)


@pytest.mark.parametrize("kind", ["boxddp", "boxfddp"])
@pytest.mark.parametrize("name", ["unicycle", "lqr_small", "lqr_large", "arm"])
def test_solve_box_same_sequence_as_oracle(name, kind):
    problem, solver, ob, xs, us, lb, ub, reginit, maxiter = _setup(name, kind)
    B = xs.shape[0]
    solver.enableTrace(maxiter + 1)
    conv = solver.solve(xs, us, maxiter, False, reginit)
    ob.solve(xs, us, maxiter, False, reginit)
    so, sg = ob.scalars(), solver.scalars()
    same = so["n_records"] == sg["n_records"]
    # solves of tens of iterations with bang-bang controls (the arm case) amplify last-bit differences until an accept
    # test flips, like the long unicycle solves (profiles/r02_sincos_sensitivity.json); the short ones must agree
    assert same.mean() >= (0.5 if name == "arm" else 0.9), (so["n_records"], sg["n_records"])
    assert np.all(same[so["n_records"] <= 20]), (so["n_records"], sg["n_records"])
    assert np.array_equal(so["converged"].astype(bool)[same], np.atleast_1d(conv)[same])
    np.testing.assert_allclose(sg["cost"][same], so["cost"][same], rtol=1e-8)
    U = solver.get("us")
    assert np.all(U >= lb - 1e-15) and np.all(U <= ub + 1e-15)
    assert ((U == lb) | (U == ub)).any(), "no active bound at the solution"
    np.testing.assert_allclose(U[same], ob.get_all("us")[same], rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(solver.get("xs")[same], ob.get_all("xs")[same], rtol=1e-6, atol=1e-6)
    for b in np.where(same)[0][:8]:
        tg, to = solver.trace(int(b)), ob.trace(int(b))
        # iterations whose expected improvement is at rounding level (dVexp below 1e-9 of the cost: the last steps under
        # the loose th_stop = 5e-5 of the box solvers) decide their step length on noise; the others must be identical
        sig = np.abs(to[:, 10]) > 1e-9 * np.abs(to[:, 1])
        k = len(sig) if sig.all() else int(np.argmin(sig))
        for col in (0, 5, 7, 8, 11):  # iter, xreg, steplength, feasibility, accepted step index
            assert np.array_equal(tg[:k, col], to[:k, col]), (b, col, tg[:, col], to[:, col])


def test_box_solver_without_limits_is_the_plain_solver():
    mod, x0 = _lqr(4, 6, 3, 7)
    T = 10
    xs, us = np.repeat(x0[:, None, :], T + 1, 1), np.zeros((4, T, 3))
    res = []
    for S in (cb.SolverBoxFDDP, cb.SolverFDDP):
        s = S(cb.ShootingProblem(x0, [mod] * T, mod))
        s.th_stop = 5e-5
        s.solve(xs, us, 20, False, 1e-9)
        res.append((s.xs, s.us, s.scalars()["n_records"]))
    assert np.array_equal(res[0][2], res[1][2])
    np.testing.assert_array_equal(res[0][0], res[1][0])
    np.testing.assert_array_equal(res[0][1], res[1][1])


def test_control_limited_lqr_reaches_scipy_optimum():
    """The same third-party pin as the oracle's (tests/test_box_oracle_cpu.py), for the CUDA path directly."""
    from scipy.optimize import minimize
    from test_box_oracle_cpu import _cost_and_gradient, _lqr_case
    P, x0, n, m, T, B, M = _lqr_case(3, fu=0.3)
    lim = 0.3
    mod = rc_lqr(n, m, P)
    problem = cb.ShootingProblem(x0, [mod] * T, mod)
    problem.set_u_bounds(-lim * np.ones(m), lim * np.ones(m))
    s = cb.SolverBoxFDDP(problem)
    conv = s.solve(np.repeat(x0[:, None, :], T + 1, 1), np.zeros((B, T, m)), 100, False, 1e-9)
    assert np.all(conv)
    U = s.us
    for b in range(B):
        ref = minimize(_cost_and_gradient, np.zeros(T * m), args=(M[b], x0[b], T, m), jac=True, method="L-BFGS-B",
                       bounds=[(-lim, lim)] * (T * m), options=dict(maxiter=5000, ftol=1e-16, gtol=1e-12))
        np.testing.assert_allclose(s.scalars()["cost"][b], ref.fun, rtol=1e-9)
        np.testing.assert_allclose(U[b].ravel(), ref.x, atol=2e-5)


def rc_lqr(n, m, P):
    from golden.ref_checks import _RawLqr
    mod = _RawLqr(n, m, P)
    mod.u_lb = mod.u_ub = None
    return mod
