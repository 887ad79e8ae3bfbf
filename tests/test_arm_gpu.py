"""GPU parity tests of the arm path (IntegratedActionModelEuler o DifferentialActionModelFreeFwdDynamics on a fixed-base
revolute chain; BASELINE configs C4/C5) against the CPU oracle, through the C ABI. Patterns of the reference's tests:
unittest/bindings/test_actions.py / test_shooting.py (calc, calcDiff, rollout of the C++ model vs a second statement),
unittest/test_diff_actions.cpp:67-84 (quasiStatic => zero acceleration), unittest/bindings/test_solvers.py:99-144
(FDDP on the Talos arm: step-wise Q*/V*/K/k, tryStep, expectedImprovement, solve). Tolerance 1e-9 relative."""
import numpy as np
import pytest

import crocoddyl_b200 as cb
import oracle_lib as ol
from test_oracle_arm_cpu import _random_chain

pytestmark = pytest.mark.gpu


def _model(kind, nv, seed=0, **kw):
    chain = cb.synthetic_arm(nv) if kind == "synthetic" else _random_chain(nv, seed)
    return cb.ArmActionModel(chain, 1e-3, **kw)


def _pair(solver_kind, model, x0, T):
    problem = cb.ShootingProblem(x0, [model] * T, model.terminal_model)
    solver = (cb.SolverFDDP if solver_kind == "fddp" else cb.SolverDDP)(problem)
    ob = ol.OracleBatch("arm", solver_kind, T, model.nx, model.nu, model.params(), np.atleast_2d(x0))
    return problem, solver, ob


def _close(a, b, name, rtol=1e-9):
    scale = max(1.0, float(np.max(np.abs(b))))
    np.testing.assert_allclose(a, b, rtol=rtol, atol=rtol * scale, err_msg=name)


def _c4_inputs(model, B, T, seed=2):
    """q ~ U(-1,1)^nv, v = 0, init_xs = {x0}, init_us = quasiStatic(x0)."""
    nv = model.nv
    rng = np.random.default_rng(seed)
    x0 = np.concatenate([rng.uniform(-1, 1, (B, nv)), np.zeros((B, nv))], 1)
    return x0, np.repeat(x0[:, None, :], T + 1, 1)


@pytest.mark.parametrize("kind,nv", [("synthetic", 7), ("random", 7), ("random", 4), ("random", 2)])
def test_problem_calc_calcdiff_rollout_quasistatic(kind, nv):
    model = _model(kind, nv, seed=4, w_xReg=1e-2, w_uReg=1e-3)
    n, T, B = 2 * nv, 6, 5
    rng = np.random.default_rng(nv)
    x0 = np.concatenate([rng.uniform(-1.5, 1.5, (B, nv)), rng.standard_normal((B, nv))], 1)
    problem, solver, ob = _pair("fddp", model, x0, T)
    xs = np.concatenate([rng.uniform(-1.5, 1.5, (B, T + 1, nv)), rng.standard_normal((B, T + 1, nv))], 2)
    us = 3. * rng.standard_normal((B, T, nv))
    ob.set_candidate(xs, us, False)
    _close(problem.calc(xs, us), ob.calc(), "total cost")
    problem.calcDiff(xs, us)
    ob.calcdiff()
    for b in (0, B - 1):
        for t in (0, T - 1, T):
            for f in ("xnext", "Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu"):
                if t == T and f in ("Lu", "Luu"):
                    continue
                _close(problem.data(f, b, t), ob.get(f, b, t), "%s b=%d t=%d" % (f, b, t))
            _close(problem.data("cost", b, t), ob.get("node_cost", b, t), "node cost")
    # quasiStatic: gravity torques; forward dynamics there gives zero acceleration (test_diff_actions.cpp:67-84)
    xq = x0.copy()
    xq[:, nv:] = 0.
    uq = problem.quasiStatic(xq)
    for b in range(B):
        _close(uq[b], ol.arm_op(model.params(), "quasistatic", xq[b, :nv], nv), "quasiStatic")
        a = ol.arm_op(model.params(), "aba", np.concatenate([xq[b], uq[b]]), nv)
        assert np.max(np.abs(a)) <= 1e-8
    # rollout (shooting.hxx:198-223): xs[t+1] = xnext(xs[t], us[t])
    xr = problem.rollout(us)
    np.testing.assert_allclose(xr[:, 0], x0)
    ob.set_candidate(xr, us, True)
    ob.calc()
    for b in (0, B - 1):
        for t in range(T):
            _close(xr[b, t + 1], ob.get("xnext", b, t), "rollout")


@pytest.mark.parametrize("solver_kind", ["ddp", "fddp"])
@pytest.mark.parametrize("kind,nv,T", [("synthetic", 7, 12), ("random", 4, 9), ("random", 7, 5)])
@pytest.mark.parametrize("feasible", [False, True])
def test_stepwise_against_oracle(solver_kind, kind, nv, T, feasible):
    model = _model(kind, nv, seed=T, w_xReg=1e-3, w_uReg=1e-3)
    B = 4
    rng = np.random.default_rng(100 + T)
    x0 = np.concatenate([rng.uniform(-1, 1, (B, nv)), 0.5 * rng.standard_normal((B, nv))], 1)
    problem, solver, ob = _pair(solver_kind, model, x0, T)
    us = rng.standard_normal((B, T, nv))
    if feasible:
        xs = problem.rollout(us)
    else:
        xs = np.concatenate([rng.uniform(-1, 1, (B, T + 1, nv)), 0.5 * rng.standard_normal((B, T + 1, nv))], 2)
    solver.setCandidate(xs, us, feasible)
    ob.set_candidate(xs, us, feasible)
    solver.x_reg = solver.u_reg = 1e-2
    ob.set("xreg", 1e-2)
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


def _compare_solve(solver, ob, conv, rtol=1e-9):
    so, sg = ob.scalars(), solver.scalars()
    assert np.array_equal(so["n_records"], sg["n_records"]), (so["n_records"], sg["n_records"])
    assert np.array_equal(so["converged"].astype(bool), np.atleast_1d(conv))
    assert np.array_equal(so["iter"], sg["iter"])
    assert np.array_equal(so["is_feasible"], sg["is_feasible"])
    for k in ("xreg", "ureg", "steplength"):
        assert np.array_equal(so[k], sg[k]), k
    _close(sg["cost"], so["cost"], "cost", rtol)
    for name in ("xs", "us", "K", "k", "Vx", "Vxx"):
        _close(solver.get(name), ob.get_all(name), name, rtol)


@pytest.mark.parametrize("solver_kind", ["fddp", "ddp"])
@pytest.mark.parametrize("kind,nv,T", [("synthetic", 7, 30), ("random", 4, 20)])
def test_solve_same_iterations_and_accept_sequence(solver_kind, kind, nv, T):
    model = _model(kind, nv, seed=9)
    B = 6
    x0, xs = _c4_inputs(model, B, T)
    problem, solver, ob = _pair(solver_kind, model, x0, T)
    us = np.repeat(problem.quasiStatic(x0)[:, None, :], T, 1)
    solver.enableTrace(100)
    conv = solver.solve(xs, us, 100, False, 0.1)  # reginit = 0.1 as benchmark/arm-manipulation-optctrl.cpp:61
    ob.solve(xs, us, 100, False, 0.1)
    # the iterates agree to ~1e-12 per iteration; over tens of iterations the comparison is made at 1e-7 on the
    # trajectories (the discrete decisions below are compared exactly)
    _compare_solve(solver, ob, conv, rtol=1e-7)
    for b in range(B):
        tg, to = solver.trace(b), ob.trace(b)
        assert tg.shape == to.shape
        for col in (0, 5, 6, 7, 8, 11, 12):  # iter, xreg, ureg, steplength, is_feasible, accepted alpha, retries
            assert np.array_equal(tg[:, col], to[:, col]), (b, col)
        np.testing.assert_allclose(tg[:, 1], to[:, 1], rtol=1e-9)  # cost per iteration


@pytest.mark.parametrize("solver_kind", ["fddp", "ddp"])
@pytest.mark.parametrize("wide", ["0", "1"])
def test_thread_per_problem_rollout_of_large_batches(solver_kind, wide, monkeypatch):
    """Batches of >= 2048 arm problems roll out with one thread per problem (cta_forward_thread) instead of one warp:
    forced here on a small batch (CGX_CTA_TPP_FORWARD_MIN), sequential and concurrent line search. Same decisions as
    the oracle in every iteration; the single-thread node function sums its 6x6 products in another order than the
    warp version, so values are compared at the usual tolerances, not bitwise with the warp schedule."""
    monkeypatch.setenv("CGX_CTA_TPP_FORWARD_MIN", "1")
    monkeypatch.setenv("CGX_CTA_WIDE", wide)
    model = _model("synthetic", 7, seed=9)
    B, T = 37, 30
    x0, xs = _c4_inputs(model, B, T)
    problem, solver, ob = _pair(solver_kind, model, x0, T)
    us = np.repeat(problem.quasiStatic(x0)[:, None, :], T, 1)
    solver.enableTrace(100)
    conv = solver.solve(xs, us, 100, False, 0.1)
    ob.solve(xs, us, 100, False, 0.1)
    _compare_solve(solver, ob, conv, rtol=1e-7)
    for b in range(0, B, 5):
        tg, to = solver.trace(b), ob.trace(b)
        assert tg.shape == to.shape
        for col in (0, 5, 6, 7, 8, 11, 12):
            assert np.array_equal(tg[:, col], to[:, col]), (b, col)
        np.testing.assert_allclose(tg[:, 1], to[:, 1], rtol=1e-9)
    # node data left by the accepted rollout (datas[t]->xnext / cost) are those of the final trajectory
    xsg, usg = solver.get("xs"), solver.get("us")
    cost = problem.calc(xsg, usg)
    np.testing.assert_allclose(cost, solver.scalars()["cost"], rtol=1e-9)


def test_config_c4_full_horizon_sampled_against_oracle():
    """BASELINE config C4 shape (7-DoF synthetic arm, T = 100, FDDP, reginit 0.1) on a batch, a sample of which is
    solved by the oracle; plus size-independent properties on the whole batch."""
    model = _model("synthetic", 7)
    B, T, NS = 128, 100, 6
    x0, xs = _c4_inputs(model, B, T)
    problem = cb.ShootingProblem(x0, [model] * T, model.terminal_model)
    solver = cb.SolverFDDP(problem)
    solver.setStorage(False)
    us = np.repeat(problem.quasiStatic(x0)[:, None, :], T, 1)
    conv = solver.solve(xs, us, 100, False, 0.1)
    sg = solver.scalars()
    assert conv.mean() > 0.9
    c = conv.astype(bool)
    assert np.all(sg["is_feasible"][c] == 1) and np.all(sg["stop"][c] < 1e-9)
    xg, ug = solver.get("xs"), solver.get("us")
    np.testing.assert_allclose(xg[:, 0], x0, atol=1e-12)  # feasible => xs[0] = x0 and gaps closed
    xr = problem.rollout(ug)
    np.testing.assert_allclose(xr[c], xg[c], rtol=1e-9, atol=1e-9)
    cost0 = problem.calc(xs, us)
    assert np.all(sg["cost"][c] < cost0[c])
    ob = ol.OracleBatch("arm", "fddp", T, 14, 7, model.params(), x0[:NS], fast=False)
    ob.solve(xs[:NS], us[:NS], 100, False, 0.1)
    so = ob.scalars()
    assert np.array_equal(so["n_records"], sg["n_records"][:NS])
    _close(sg["cost"][:NS], so["cost"], "cost")
    _close(xg[:NS], ob.get_all("xs"), "xs", 1e-7)
    _close(ug[:NS], ob.get_all("us"), "us", 1e-7)


def _solve_in_mode(env, solver_kind, model, x0, xs, T, maxiter=100):
    """One batched solve with the development knobs of the CTA backend set (read when the solver is created)."""
    import os
    keys = ("CGX_CTA_FUSED", "CGX_CTA_PHASES", "CGX_CTA_WIDE")
    old = {k: os.environ.pop(k, None) for k in keys}
    os.environ.update(env)
    try:
        problem = cb.ShootingProblem(x0, [model] * T, model.terminal_model)
        us = np.repeat(problem.quasiStatic(x0)[:, None, :], T, 1)
        solver = (cb.SolverFDDP if solver_kind == "fddp" else cb.SolverDDP)(problem)
        solver.enableTrace(maxiter + 1)
        solver.solve(xs, us, maxiter, False, 0.1)
        out = {"sc": solver.scalars(), "tr": [solver.trace(b) for b in range(x0.shape[0])]}
        for f in ("xs", "us", "K", "k", "Vx", "Vxx"):
            out[f] = solver.get(f)
        return out
    finally:
        for k in keys:
            os.environ.pop(k, None)
            if old[k] is not None:
                os.environ[k] = old[k]


@pytest.mark.parametrize("solver_kind", ["fddp", "ddp"])
def test_phase_split_and_concurrent_line_search_equal_the_fused_loop(solver_kind):
    """The three schedules of the same solve loop — fused per-problem kernel, lockstep phase launches with a
    sequential line search, and phase launches with 4 step lengths rolled out at once — must make identical
    decisions (iterations, accepted step length, regularisation, counters) and leave the same iterates."""
    model = _model("synthetic", 7)
    B, T = 48, 25
    x0, xs = _c4_inputs(model, B, T, seed=3)
    ref = _solve_in_mode({"CGX_CTA_FUSED": "1"}, solver_kind, model, x0, xs, T)
    assert (ref["sc"]["n_forward"] > ref["sc"]["n_backward"]).any()  # the case does reject step lengths
    for env in ({"CGX_CTA_WIDE": "0"}, {"CGX_CTA_WIDE": "1"}, {}):
        got = _solve_in_mode(env, solver_kind, model, x0, xs, T)
        for k in ("iter", "converged", "n_records", "n_backward", "n_forward", "is_feasible", "xreg", "steplength"):
            assert np.array_equal(got["sc"][k], ref["sc"][k]), (env, k)
        for b in range(B):
            assert got["tr"][b].shape == ref["tr"][b].shape, (env, b)
            for col in (0, 5, 6, 7, 8, 11, 12):
                assert np.array_equal(got["tr"][b][:, col], ref["tr"][b][:, col]), (env, b, col)
        _close(got["sc"]["cost"], ref["sc"]["cost"], "cost", 1e-12)
        for f in ("xs", "us", "K", "k", "Vx", "Vxx"):
            _close(got[f], ref[f], f, 1e-10)


@pytest.mark.parametrize("alphas", [[1.0], [1.0, 0.5], [1.0, 0.5, 0.25, 0.125, 0.0625], [1.0, 0.7, 0.3, 0.1, 0.05, 0.01, 1e-3]])
def test_concurrent_line_search_with_other_step_length_lists(alphas):
    """SolverDDP::set_alphas (ddp.cpp:444-462) with fewer / more step lengths than the 4 slots of the concurrent line
    search (small batches run it from the first round): same decisions and iterates as the oracle."""
    model = _model("synthetic", 7)
    B, T = 10, 20
    x0, xs = _c4_inputs(model, B, T, seed=5)
    problem, solver, ob = _pair("fddp", model, x0, T)
    us = np.repeat(problem.quasiStatic(x0)[:, None, :], T, 1)
    solver.alphas = alphas
    ob.set_alphas(alphas)
    solver.enableTrace(40)
    conv = solver.solve(xs, us, 30, False, 0.1)
    ob.solve(xs, us, 30, False, 0.1)
    _compare_solve(solver, ob, conv, rtol=1e-7)
    sg, so = solver.scalars(), ob.scalars()
    assert np.array_equal(sg["steplength"], so["steplength"])
    for b in range(B):
        tg, to = solver.trace(b), ob.trace(b)
        assert tg.shape == to.shape
        for col in (0, 5, 6, 7, 8, 11, 12):
            assert np.array_equal(tg[:, col], to[:, col]), (b, col)


@pytest.mark.parametrize("maxiter,feasible", [(0, False), (1, False), (3, True)])
def test_phase_split_solve_small_iteration_budgets(maxiter, feasible):
    """maxiter = 0 returns the candidate untouched, maxiter = 1 is the reference benchmark's call
    (benchmark/arm-manipulation-optctrl.cpp:61), and a feasible (rolled-out) start takes the closed-gap branches."""
    model = _model("synthetic", 7)
    B, T = 5, 15
    x0, xs = _c4_inputs(model, B, T, seed=8)
    problem, solver, ob = _pair("fddp", model, x0, T)
    us = np.repeat(problem.quasiStatic(x0)[:, None, :], T, 1)
    if feasible:
        xs = problem.rollout(us)
    conv = solver.solve(xs, us, maxiter, feasible, 0.1)
    ob.solve(xs, us, maxiter, feasible, 0.1)
    sg, so = solver.scalars(), ob.scalars()
    assert np.array_equal(sg["iter"], so["iter"]) and np.array_equal(sg["n_records"], so["n_records"])
    assert np.array_equal(np.atleast_1d(conv), so["converged"].astype(bool))
    _close(solver.get("xs"), ob.get_all("xs"), "xs", 1e-9)
    _close(solver.get("us"), ob.get_all("us"), "us", 1e-9)
    if maxiter:
        _close(sg["cost"], so["cost"], "cost", 1e-9)


def test_config_c5_full_size_properties():
    """BASELINE config C5 at its full size (8192 arm OCPs, T = 100) through size-independent properties: every problem
    converges to a feasible trajectory that reproduces under rollout and lowers the cost; the result of a problem does
    not depend on the batch it was solved in (a strided sample re-solved as a small batch — which runs the concurrent
    line search from the first round, the full batch only in its tail — makes the same decisions); a sample agrees
    with the oracle; and re-solving from the solution converges again without raising the cost."""
    model = _model("synthetic", 7)
    B, T = 8192, 100
    x0, xs = _c4_inputs(model, B, T)
    problem = cb.ShootingProblem(x0, [model] * T, model.terminal_model)
    solver = cb.SolverFDDP(problem)
    solver.setStorage(False)
    us = np.repeat(problem.quasiStatic(x0)[:, None, :], T, 1)
    conv = solver.solve(xs, us, 100, False, 0.1)
    sg = solver.scalars()
    assert conv.all()
    assert np.all(sg["is_feasible"] == 1) and np.all(sg["stop"] < 1e-9) and np.all(np.isfinite(sg["cost"]))
    xg, ug = solver.get("xs"), solver.get("us")
    np.testing.assert_allclose(xg[:, 0], x0, atol=1e-12)
    np.testing.assert_allclose(problem.rollout(ug), xg, rtol=1e-9, atol=1e-9)
    assert np.all(sg["cost"] < problem.calc(xs, us))
    # batch independence on a strided sample that includes the slowest problems
    idx = np.unique(np.concatenate([np.arange(0, B, 171), np.argsort(sg["n_forward"])[-8:]]))
    p2 = cb.ShootingProblem(x0[idx], [model] * T, model.terminal_model)
    s2 = cb.SolverFDDP(p2)
    s2.setStorage(False)
    s2.solve(xs[idx], us[idx], 100, False, 0.1)
    g2 = s2.scalars()
    for k in ("iter", "n_records", "n_backward", "n_forward", "xreg", "steplength", "converged"):
        assert np.array_equal(g2[k], sg[k][idx]), k
    _close(g2["cost"], sg["cost"][idx], "cost", 1e-12)
    _close(s2.get("xs"), xg[idx], "xs", 1e-10)
    _close(s2.get("us"), ug[idx], "us", 1e-10)
    # a sample against the oracle
    NS = 4
    ob = ol.OracleBatch("arm", "fddp", T, 14, 7, model.params(), x0[idx[:NS]], fast=False)
    ob.solve(xs[idx[:NS]], us[idx[:NS]], 100, False, 0.1)
    so = ob.scalars()
    assert np.array_equal(so["n_records"], g2["n_records"][:NS])
    _close(g2["cost"][:NS], so["cost"], "cost")
    # re-solve from the solution (feasible start, fresh regularisation): the stopping test of this benchmark setting
    # (costs scaled by dt = 1e-3) fires before the optimum, so the second solve may still lower the cost — it must
    # converge again, stay feasible and never raise it
    s2.solve(xg[idx], ug[idx], 100, True, 0.1)
    g3 = s2.scalars()
    assert np.all(g3["converged"] == 1) and np.all(g3["is_feasible"] == 1)
    assert np.all(g3["cost"] <= g2["cost"] + 1e-9)


# ------------------------------------------------------------------ options of the differential model
_F3_OPTIONS = {
    "armature": dict(armature=[0.1, 0.2, 0.05, 0.3, 0.15, 0.02, 0.07]),  # examples/arm_manipulation.py:53
    "weighted_state": dict(xWeights=np.concatenate([np.linspace(0.5, 3., 7), np.linspace(2., 0.1, 7)]), w_xReg=1e-2),
    "frame_translation": dict(goal="translation"),
    "frame_rotation": dict(goal="rotation"),
    "all": dict(armature=[0.1] * 7, xWeights=np.linspace(0.2, 2., 14), w_xReg=1e-3, goal="translation"),
}


@pytest.mark.parametrize("name", sorted(_F3_OPTIONS))
def test_arm_model_options_node_data_and_solve(name):
    """Armature branch (free-fwddyn.hxx:66-74,108-114), ActivationModelWeightedQuad on the state cost, CostModelFrame
    Translation / Rotation as the goal cost: ShootingProblem::calc / calcDiff node data, the quasi-static torques and a
    whole FDDP solve (thread and warp rollouts) against the oracle, which takes the reference's Minv route for the
    armature while the kernels add it to the pivots of the articulated-body sweep."""
    model = _model("synthetic", 7, **_F3_OPTIONS[name])
    nv, n, T, B = 7, 14, 20, 6
    rng = np.random.default_rng(5)
    x0, xs0 = _c4_inputs(model, B, T)
    problem, solver, ob = _pair("fddp", model, x0, T)
    xs = np.concatenate([rng.uniform(-1.5, 1.5, (B, T + 1, nv)), rng.standard_normal((B, T + 1, nv))], 2)
    us = 3. * rng.standard_normal((B, T, nv))
    ob.set_candidate(xs, us, False)
    _close(problem.calc(xs, us), ob.calc(), "total cost")
    problem.calcDiff(xs, us)
    ob.calcdiff()
    for b in (0, B - 1):
        for t in (0, T - 1, T):
            for f in ("xnext", "Fx", "Fu", "Lx", "Lu", "Lxx", "Luu"):
                if t == T and f in ("Fu", "Lu", "Luu"):
                    continue
                _close(problem.data(f, b, t), ob.get(f, b, t), "%s b=%d t=%d" % (f, b, t))
            _close(problem.data("cost", b, t), ob.get("node_cost", b, t), "node cost")
    xr = problem.rollout(us)  # shooting.hxx:198-223 through the rollout kernel (warp-cooperative calc)
    ob.set_candidate(xr, us, True)
    ob.calc()
    for b in (0, B - 1):
        for t in range(T):
            _close(xr[b, t + 1], ob.get("xnext", b, t), "rollout")
    us_q = np.repeat(problem.quasiStatic(x0)[:, None, :], T, 1)
    solver.enableTrace(100)
    conv = solver.solve(xs0, us_q, 100, False, 0.1)
    ob.solve(xs0, us_q, 100, False, 0.1)
    _compare_solve(solver, ob, conv, rtol=1e-7)
    for b in range(B):
        tg, to = solver.trace(b), ob.trace(b)
        assert tg.shape == to.shape
        for col in (0, 5, 6, 7, 8, 11, 12):
            assert np.array_equal(tg[:, col], to[:, col]), (b, col)
