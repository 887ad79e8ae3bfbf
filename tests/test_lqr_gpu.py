"""GPU parity tests of the CTA-per-problem path (ActionModelLQR) against the CPU oracle, through the C ABI.
Patterns of the reference's tests: unittest/bindings/test_solvers.py:38-96 (step-wise Q*/V*/K/k), unittest/
test_solvers.cpp:65-110 (DDP/FDDP solve vs KKT on LQR incl. drift), benchmark/lqr-optctrl.cpp:16-54 (config C1).
Tolerance 1e-9 relative (BASELINE.json north_star)."""
import numpy as np
import pytest

import crocoddyl_b200 as cb
import oracle_lib as ol

pytestmark = pytest.mark.gpu


def _random_lqr(n, m, B, seed, drift_free=True):
    """Fx = I + 0.1 G/sqrt(n), Fu = G, Lxx = A'A/n + I, Luu = B'B/m + I, Lxu = 0.1 G."""
    rng = np.random.default_rng(seed)
    mod = cb.ActionModelLQR(n, m, drift_free)
    mod.Fx = np.eye(n) + 0.1 * rng.standard_normal((B, n, n)) / np.sqrt(n)
    mod.Fu = rng.standard_normal((B, n, m))
    A = rng.standard_normal((B, n, n))
    Bm = rng.standard_normal((B, m, m))
    mod.Lxx = np.einsum("bki,bkj->bij", A, A) / n + np.eye(n)
    mod.Luu = np.einsum("bki,bkj->bij", Bm, Bm) / m + np.eye(m)
    mod.Lxu = 0.1 * rng.standard_normal((B, n, m))
    mod.lx = rng.standard_normal((B, n))
    mod.lu = rng.standard_normal((B, m))
    mod.f0 = rng.standard_normal((B, n))
    x0 = rng.standard_normal((B, n))
    return mod, x0


def _pair(kind, mod, x0, T):
    problem = cb.ShootingProblem(x0, [mod] * T, mod)
    solver = (cb.SolverFDDP if kind == "fddp" else cb.SolverDDP)(problem)
    ob = ol.OracleBatch("lqr", kind, T, mod.nx, mod.nu, mod.params(), np.atleast_2d(x0))
    return problem, solver, ob


def _close(a, b, name, rtol=1e-9):
    scale = max(1.0, float(np.max(np.abs(b))))
    np.testing.assert_allclose(a, b, rtol=rtol, atol=rtol * scale, err_msg=name)


@pytest.mark.parametrize("kind", ["ddp", "fddp"])
@pytest.mark.parametrize("n,m,T", [(5, 2, 7), (12, 4, 20), (37, 12, 10), (8, 8, 5)])
@pytest.mark.parametrize("feasible", [False, True])
def test_stepwise_against_oracle(kind, n, m, T, feasible):
    B = 5
    mod, x0 = _random_lqr(n, m, B, 7 * n + T, drift_free=(n != 12))
    problem, solver, ob = _pair(kind, mod, x0, T)
    rng = np.random.default_rng(T)
    us = 0.3 * rng.standard_normal((B, T, m))
    xs = problem.rollout(us) if feasible else rng.standard_normal((B, T + 1, n))
    solver.setCandidate(xs, us, feasible)
    ob.set_candidate(xs, us, feasible)
    solver.x_reg = 1e-3
    solver.u_reg = 1e-2
    ob.set("xreg", 1e-3)
    ob.set("ureg", 1e-2)
    solver.computeDirection(True)
    ob.calcdiff()
    assert not ob.backward().any()
    for name in ("fs", "Vxx", "Vx", "K", "k", "Qxx", "Qxu", "Quu", "Qx", "Qu"):
        _close(solver.get(name), ob.get_all(name), name)
    d_o, stop_o = ob.expected_improvement()
    _close(solver.stoppingCriteria(), stop_o, "stop")
    for alpha in (1.0, 0.5):
        dV = solver.tryStep(alpha)
        dV_o, st = ob.try_step(alpha)
        assert not st.any()
        cost_scale = max(1.0, float(np.max(np.abs(ob.scalars()["cost"]))))
        np.testing.assert_allclose(dV, dV_o, rtol=1e-9, atol=1e-9 * cost_scale)
        _close(solver.get("xs_try"), ob.get_all("xs_try"), "xs_try")
        _close(solver.get("us_try"), ob.get_all("us_try"), "us_try")
        d_o, _ = ob.expected_improvement()
        _close(solver.expectedImprovement(), d_o, "d")


def _compare_solve(solver, ob, conv):
    so, sg = ob.scalars(), solver.scalars()
    assert np.array_equal(so["n_records"], sg["n_records"]), (so["n_records"], sg["n_records"])
    assert np.array_equal(so["converged"].astype(bool), np.atleast_1d(conv))
    assert np.array_equal(so["iter"], sg["iter"])
    assert np.array_equal(so["is_feasible"], sg["is_feasible"])
    for k in ("xreg", "ureg", "steplength"):
        assert np.array_equal(so[k], sg[k]), k
    _close(sg["cost"], so["cost"], "cost")
    for name in ("xs", "us", "K", "k", "Vx", "Vxx"):
        _close(solver.get(name), ob.get_all(name), name)


@pytest.mark.parametrize("kind", ["ddp", "fddp"])
@pytest.mark.parametrize("n,m,T,drift_free", [(12, 4, 30, True), (20, 10, 10, False), (36, 12, 25, True)])
def test_solve_against_oracle(kind, n, m, T, drift_free):
    B = 6
    mod, x0 = _random_lqr(n, m, B, n + T, drift_free)
    problem, solver, ob = _pair(kind, mod, x0, T)
    xs = np.repeat(x0[:, None, :], T + 1, 1)
    us = np.zeros((B, T, m))
    solver.enableTrace(100)
    conv = solver.solve(xs, us, 100, False, 1e-9)
    ob.solve(xs, us, 100, False, 1e-9)
    _compare_solve(solver, ob, conv)
    for b in range(B):
        tg, to = solver.trace(b), ob.trace(b)
        assert tg.shape == to.shape
        for col in (0, 5, 6, 7, 8, 11, 12):
            assert np.array_equal(tg[:, col], to[:, col]), (b, col)
    # LQR: DDP/FDDP = one Newton step, alpha = 1 at the first iteration (test_solvers.cpp idea)
    assert np.all([solver.trace(b)[0, 11] == 0 for b in range(B)])


def test_config_c1_lqr_optctrl_benchmark_problem():
    """BASELINE config C1 = benchmark/lqr-optctrl.cpp:16-47: ActionModelLQR(37, 12) defaults, T = 100, x0 = 0,
    xs = {x0}, us = {0}, SolverDDP::solve(xs, us, MAXITER = 1)."""
    n, m, T = 37, 12, 100
    mod = cb.ActionModelLQR(n, m, True)
    x0 = np.zeros(n)
    problem, solver, ob = _pair("ddp", mod, x0, T)
    xs = np.zeros((1, T + 1, n))
    us = np.zeros((1, T, m))
    conv = solver.solve(xs, us, 1)
    ob.solve(xs, us, 1, False, 1e-9)
    _compare_solve(solver, ob, conv)
    # the KKT solver of the reference's own test solves the same LQR problem in one step
    conv = solver.solve(xs, us, 100)
    ok, kx, ku = ob.kkt_solve(0, xs[0], us[0], 100)
    assert ok == 1
    np.testing.assert_allclose(solver.get("xs")[0], kx, atol=1e-8)
    np.testing.assert_allclose(solver.get("us")[0], ku, atol=1e-8)


def test_problem_calc_calcdiff_rollout_lqr():
    n, m, T, B = 9, 3, 6, 4
    mod, x0 = _random_lqr(n, m, B, 3, False)
    problem, solver, ob = _pair("ddp", mod, x0, T)
    rng = np.random.default_rng(2)
    xs, us = rng.standard_normal((B, T + 1, n)), rng.standard_normal((B, T, m))
    ob.set_candidate(xs, us, False)
    _close(problem.calc(xs, us), ob.calc(), "cost")
    problem.calcDiff(xs, us)
    ob.calcdiff()
    for b in (0, B - 1):
        for t in (0, T - 1, T):
            for f in ("Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu", "xnext"):
                _close(problem.data(f, b, t), ob.get(f, b, t), f)
    xr = problem.rollout(us)
    np.testing.assert_allclose(xr[:, 0], x0)
    for t in range(T):
        nxt = np.einsum("bij,bj->bi", mod.Fx, xr[:, t]) + np.einsum("bij,bj->bi", mod.Fu, us[:, t]) + mod.f0
        np.testing.assert_allclose(xr[:, t + 1], nxt, rtol=1e-9, atol=1e-9)


def test_riccati_sweep_config_c3_sizes():
    """BASELINE config C3 (nx in {12,36,72}, nu = nx/3, T = 100): one backwardPass with derivatives resident,
    checked on a small batch against the oracle's K, k, Vx, Vxx."""
    for n in (12, 36, 72):
        m, T, B = n // 3, 100, 3
        mod, x0 = _random_lqr(n, m, B, n, True)
        problem, solver, ob = _pair("fddp", mod, x0, T)
        xs = np.repeat(x0[:, None, :], T + 1, 1)
        us = np.zeros((B, T, m))
        solver.setStorage(False)
        solver.setCandidate(xs, us, False)
        ob.set_candidate(xs, us, False)
        solver.x_reg = solver.u_reg = 1e-9
        ob.set("xreg", 1e-9)
        ob.set("ureg", 1e-9)
        solver.computeDirection(True)
        solver.computeDirection(False)
        ob.calcdiff()
        assert not ob.backward().any()
        for name in ("K", "k", "Vx", "Vxx"):
            _close(solver.get(name), ob.get_all(name), "%s n=%d" % (name, n))


# ------------------------------------------------------------------ DifferentialActionModelLQR on the LQR kernels
@pytest.mark.parametrize("kind", ["ddp", "fddp"])
@pytest.mark.parametrize("drift_free", [True, False])
def test_integrated_euler_of_differential_lqr_against_the_two_step_oracle(kind, drift_free):
    """IntegratedActionModelEuler(DifferentialActionModelLQR(nq, nu), dt) with a dt = 0 terminal model, solved by the
    CUDA path (discrete matrices of the Euler step on the LQR kernels, crocoddyl_b200/integrated.py), against the
    oracle's statement of the reference's two-step evaluation (diff-lqr.hxx:48-84 then euler.hxx:65-130): node
    quantities through the plugin calls, then the whole solve."""
    from integrated_cases import oracle_params, random_difflqr
    nq, nu, B, T, dt = 6, 3, 5, 25, 0.02
    d, x0 = random_difflqr(nq, nu, B, 21, drift_free)
    run, term = cb.IntegratedActionModelEuler(d, dt), cb.IntegratedActionModelEuler(d, 0.)
    problem = cb.ShootingProblem(x0, [run] * T, term)
    solver = (cb.SolverFDDP if kind == "fddp" else cb.SolverDDP)(problem)
    ob = ol.OracleBatch("euler_difflqr", kind, T, 2 * nq, nu, oracle_params(d, dt), x0, tparams=oracle_params(d, 0.))
    rng = np.random.default_rng(3)
    xs_r, us_r = rng.standard_normal((B, T + 1, 2 * nq)), rng.standard_normal((B, T, nu))
    # ShootingProblem::calc / calcDiff on a random trajectory
    ob.set_candidate(xs_r, us_r, False)
    ob.calc()
    ob.calcdiff()
    costs = problem.calc(xs_r, us_r)
    _close(costs, np.array([sum(ob.get("node_cost", b, t)[0] for t in range(T + 1)) for b in range(B)]), "cost")
    problem.calcDiff(xs_r, us_r)
    for name in ("Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu"):
        for b in (0, B - 1):
            for t in (0, T // 2, T):
                if t == T and name in ("Fu", "Lu", "Lxu", "Luu"):
                    continue  # control derivatives of the terminal node are not used (action-base.hxx:27-37)
                _close(problem.data(name, b, t), ob.get(name, b, t), "%s b=%d t=%d" % (name, b, t))
    xs = np.repeat(x0[:, None, :], T + 1, 1)
    us = np.zeros((B, T, nu))
    conv = solver.solve(xs, us, 30, False, 1e-9)
    ob.solve(xs, us, 30, False, 1e-9)
    _compare_solve(solver, ob, conv)


@pytest.mark.parametrize("name", ["driftfree", "drift"])
def test_integrated_differential_lqr_node_data_equal_the_reference_vectors(name):
    """The CUDA path (through the C ABI) against vectors the reference's legacy NumPy classes produced for
    IntegratedActionModelEuler(DifferentialActionModelLQR) (tests/golden/ref_golden_difflqr.npz): xnext, Fx, Fu as the
    reference computes them; cost and its derivatives times dt (the C++ integrator scales them, euler.hxx:68,117-121)."""
    import os
    G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_golden_difflqr.npz"))
    pre = "euler_difflqr_%s/" % name
    c = {k[len(pre):]: G[k] for k in G.files if k.startswith(pre)}
    dt, N = float(c["dt"]), len(c["x"])
    d = cb.DifferentialActionModelLQR(3, 2, bool(c["drift_free"]))
    for k in ("Fq", "Fv", "Fu", "f0", "Lxx", "Lxu", "Luu", "lx", "lu"):
        setattr(d, k, np.repeat(c["p_" + k][None], N, 0))  # one problem per sample point
    model = cb.IntegratedActionModelEuler(d, dt)
    problem = cb.ShootingProblem(c["x"], [model], model)
    xs = np.repeat(c["x"][:, None, :], 2, 1)
    us = c["u"][:, None, :]
    problem.calc(xs, us)
    problem.calcDiff(xs, us)
    tol = dict(rtol=1e-12, atol=1e-13)
    for b in range(N):
        np.testing.assert_allclose(problem.data("xnext", b, 0), c["xnext"][b], **tol)
        np.testing.assert_allclose(problem.data("Fx", b, 0), c["Fx"][b], **tol)
        np.testing.assert_allclose(problem.data("Fu", b, 0), c["Fu"][b], **tol)
        np.testing.assert_allclose(problem.data("cost", b, 0)[0], dt * c["cost"][b], **tol)
        for f in ("Lx", "Lu", "Lxx", "Lxu", "Luu"):
            np.testing.assert_allclose(problem.data(f, b, 0), dt * c[f][b], err_msg=f, **tol)
