"""The plugin half of the boundary (include/crocoddyl/core/action-base.hpp:34-142): ActionModelAbstract.calc /
calcDiff / createData / quasiStatic per node, and ShootingProblem(x0, running_models, terminal_model) with a terminal
model of its own (examples/arm_manipulation.py:41-63: terminal cost = gripperPose only). Against the reference-produced
vectors (unicycle / LQR nodes) and the CPU oracle (arm, terminal models)."""
import numpy as np
import pytest

import crocoddyl_b200 as cb
import oracle_lib as ol
from golden import ref_checks as rc

pytestmark = pytest.mark.gpu


def test_unicycle_model_calc_calcdiff_vs_reference_vectors():
    g = rc.load()["unicycle_node"]
    m = cb.ActionModelUnicycle()
    d = m.createData()
    assert m.checkData(d) and m.nu == 2 and m.nr == 5 and m.state.nx == 3
    for x, u, i in zip(g["x"], g["u"], range(len(g["x"]))):
        m.calc(d, x, u)
        rc.close(d.cost, g["cost"][i], "cost")
        rc.close(d.xnext, g["xnext"][i], "xnext")
        m.calcDiff(d, x, u)
        for f in ("Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu"):
            rc.close(getattr(d, f), g[f][i], f)
    m.calcDiff(d, g["x"][0])  # x-only overload: u = unone
    rc.close(d.cost, g["term_cost"], "terminal cost")
    rc.close(d.Lx, g["term_Lx"], "terminal Lx")
    # batched call: every field gets the batch dimension
    m.calcDiff(d, g["x"], g["u"])
    rc.close(d.Fx, g["Fx"], "batched Fx")
    rc.close(d.cost, g["cost"], "batched cost")


@pytest.mark.parametrize("cname", ["lqr_node_driftfree", "lqr_node_drift"])
def test_lqr_model_calc_calcdiff_vs_reference_vectors(cname):
    g = rc.load()[cname]
    n, m_ = g["p_Fu"].shape
    m = cb.ActionModelLQR(n, m_, bool(g["drift_free"]))
    for k in ("Fx", "Fu", "f0", "Lxx", "Lxu", "Luu", "lx", "lu"):
        setattr(m, k, g["p_" + k])
    d = m.createData()
    for i, (x, u) in enumerate(zip(g["x"], g["u"])):
        m.calcDiff(d, x, u)
        for f in ("cost", "xnext", "Lx", "Lu"):
            rc.close(getattr(d, f), g[f][i], f)
        rc.close(d.Fx, g["p_Fx"], "Fx")
        rc.close(d.Luu, g["p_Luu"], "Luu")


def test_arm_model_node_calls_and_quasistatic():
    nv = 7
    arm = cb.ArmActionModel(cb.synthetic_arm(nv), 1e-3)
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.uniform(-1, 1, nv), 0.3 * rng.standard_normal(nv)])
    u = rng.standard_normal(nv)
    d = arm.createData()
    arm.calcDiff(d, x, u)
    ob = ol.OracleBatch("arm", "ddp", 1, 2 * nv, nv, arm.params(), x[None])
    ob.set_candidate(np.stack([x, x])[None], u[None, None], False)
    ob.calc()
    ob.calcdiff()
    np.testing.assert_allclose(d.cost, ob.get("node_cost", 0, 0)[0], rtol=1e-10)
    for f in ("xnext", "Fx", "Fu", "Lx", "Lu", "Lxx", "Luu"):
        ref = ob.get(f, 0, 0)
        np.testing.assert_allclose(getattr(d, f), ref, rtol=1e-9, atol=1e-10 * max(1., np.abs(ref).max()), err_msg=f)
    uq = arm.quasiStatic(d, x)
    np.testing.assert_allclose(uq, ol.arm_op(arm.params(), "quasistatic", x[:nv], nv), rtol=1e-10, atol=1e-12)
    # quasiStatic => zero acceleration from rest (unittest/test_diff_actions.cpp:67-84)
    x_rest = np.concatenate([x[:nv], np.zeros(nv)])
    arm.calc(d, x_rest, arm.quasiStatic(d, x_rest))
    np.testing.assert_allclose(d.xnext, x_rest, atol=1e-9)


@pytest.mark.parametrize("kind", ["ddp", "fddp"])
def test_arm_terminal_model_with_its_own_cost_set(kind):
    """examples/arm_manipulation.py:41-63: running costs gripperPose + xReg + uReg, TERMINAL cost = gripperPose only,
    terminal model = Euler(DAM(terminal costs), dt = 0)."""
    nv, T, B = 7, 15, 4
    chain = cb.synthetic_arm(nv)
    run = cb.ArmActionModel(chain, 1e-3, w_gripperPose=1., w_xReg=1e-4, w_uReg=1e-4)
    term = cb.ArmActionModel(chain, 0., w_gripperPose=1., w_xReg=0., w_uReg=0.)
    rng = np.random.default_rng(3)
    x0 = np.concatenate([rng.uniform(-1, 1, (B, nv)), np.zeros((B, nv))], 1)
    problem = cb.ShootingProblem(x0, [run] * T, term)
    xs = np.repeat(x0[:, None, :], T + 1, 1)
    us = np.repeat(problem.quasiStatic(x0)[:, None, :], T, 1)
    ob = ol.OracleBatch("arm", kind, T, 2 * nv, nv, run.params(), x0, tparams=term.params())
    # terminal data differ from the default terminal (same cost set, dt = 0)
    problem.calc(xs, us)
    problem.calcDiff(xs, us)
    ob.set_candidate(xs, us, False)
    ob.calc()
    ob.calcdiff()
    for b in range(B):
        np.testing.assert_allclose(problem.data("cost", b, T)[0], ob.get("node_cost", b, T)[0], rtol=1e-10)
        np.testing.assert_allclose(problem.data("Lxx", b, T), ob.get("Lxx", b, T), rtol=1e-9, atol=1e-12)
        assert np.all(problem.data("Lxx", b, T)[nv:, nv:] == 0.)  # no xReg at the terminal node: no velocity Hessian
    solver = (cb.SolverFDDP if kind == "fddp" else cb.SolverDDP)(problem)
    solver.enableTrace(60)
    solver.solve(xs, us, 50, False, 0.1)
    ob.solve(xs, us, 50, False, 0.1)
    so, sg = ob.scalars(), solver.scalars()
    assert np.array_equal(so["n_records"], sg["n_records"])
    np.testing.assert_allclose(sg["cost"], so["cost"], rtol=1e-9)
    np.testing.assert_allclose(solver.xs, ob.get_all("xs"), rtol=1e-7, atol=1e-7)
    for b in range(B):
        assert np.array_equal(solver.trace(b)[:, (5, 7, 8, 11)], ob.trace(b)[:, (5, 7, 8, 11)])
    # and it is NOT the solution of the problem with the default terminal model
    s2 = cb.SolverFDDP(cb.ShootingProblem(x0, [run] * T, run.terminal_model))
    s2.solve(xs, us, 50, False, 0.1)
    assert np.max(np.abs(s2.scalars()["cost"] - sg["cost"])) > 1e-9


@pytest.mark.parametrize("family", ["unicycle", "lqr"])
def test_terminal_model_of_its_own_unicycle_and_lqr(family):
    T, B = 12, 5
    rng = np.random.default_rng(8)
    if family == "unicycle":
        run, term = cb.ActionModelUnicycle(), cb.ActionModelUnicycle()
        term.costWeights = np.array([30., 1.])  # heavier terminal state cost
        x0 = np.stack([rng.uniform(-2, 2, B), rng.uniform(-2, 2, B), rng.uniform(-np.pi, np.pi, B)], 1)
    else:
        n, m = 6, 3
        run, term = cb.ActionModelLQR(n, m, False), cb.ActionModelLQR(n, m, False)
        run.Fx = np.eye(n) + 0.1 * rng.standard_normal((n, n))
        run.Fu = rng.standard_normal((n, m))
        term.Lxx = 10. * np.eye(n)
        term.lx = rng.standard_normal(n)
        x0 = rng.standard_normal((B, n))
    problem = cb.ShootingProblem(x0, [run] * T, term)
    xs, us = np.repeat(x0[:, None, :], T + 1, 1), np.zeros((B, T, run.nu))
    for kind, S in (("ddp", cb.SolverDDP), ("fddp", cb.SolverFDDP)):
        solver = S(problem)
        conv = solver.solve(xs, us, 100, False, 1e-9)
        ob = ol.OracleBatch(family, kind, T, run.nx, run.nu, run.params(), x0, tparams=term.params())
        ob.solve(xs, us, 100, False, 1e-9)
        so, sg = ob.scalars(), solver.scalars()
        short = so["n_records"] <= 60
        assert np.array_equal(so["n_records"][short], sg["n_records"][short])
        np.testing.assert_allclose(sg["cost"][short], so["cost"][short], rtol=1e-9)
        np.testing.assert_allclose(solver.xs[short], ob.get_all("xs")[short], rtol=1e-8, atol=1e-8)
        np.testing.assert_allclose(solver.get("Vxx")[short][:, T], ob.get_all("Vxx")[short][:, T], rtol=1e-9, atol=1e-12)
