"""IntegratedActionModelEuler(DifferentialActionModelLQR) is run on the LQR kernels through discrete
matrices built on the host (crocoddyl_b200/integrated.py). This CPU test holds that mapping to the oracle's two-step
statement of the reference (diff-lqr.hxx:48-84 evaluated first, euler.hxx:65-74,100-130 after): node calc / calcDiff
and complete DDP / FDDP solves agree to 1e-9 when the ORACLE's ActionModelLQR is given the mapped matrices."""
import numpy as np
import pytest

import crocoddyl_b200 as cb
import oracle_lib as ol
from integrated_cases import oracle_params, random_difflqr


def _close(a, b, name, rtol=1e-9):
    scale = max(1.0, float(np.max(np.abs(b))))
    np.testing.assert_allclose(a, b, rtol=rtol, atol=rtol * scale, err_msg=name)


def test_reference_constructor_defaults():
    d = cb.DifferentialActionModelLQR(3, 2)
    assert d.driftFree and d.nx == 6 and d.nu == 2
    np.testing.assert_array_equal(d.Fq, np.eye(3))
    np.testing.assert_array_equal(d.Fu, np.eye(3, 2))
    np.testing.assert_array_equal(d.f0, np.ones(3))
    np.testing.assert_array_equal(d.Lxu, np.eye(6, 2))
    m = cb.IntegratedActionModelEuler(d)  # timeStep = 1e-3 (euler.hpp)
    assert m.timeStep == 1e-3 and m.nx == 6 and m.nu == 2
    assert cb.IntegratedActionModelEuler(d, -1.).timeStep == 1e-3  # euler.hxx:27-31
    with pytest.raises(TypeError):
        cb.IntegratedActionModelEuler(cb.ActionModelUnicycle(), 0.1)


@pytest.mark.parametrize("drift_free", [True, False])
@pytest.mark.parametrize("dt", [0.05, 0.])
def test_node_calc_and_calcdiff_of_the_mapping(drift_free, dt):
    nq, nu, B, T = 3, 2, 4, 3
    d, x0 = random_difflqr(nq, nu, B, 5, drift_free)
    m = cb.IntegratedActionModelEuler(d, dt)
    rng = np.random.default_rng(1)
    xs, us = rng.standard_normal((B, T + 1, 2 * nq)), rng.standard_normal((B, T, nu))
    a = ol.OracleBatch("lqr", "ddp", T, 2 * nq, nu, m.params(), x0)
    b = ol.OracleBatch("euler_difflqr", "ddp", T, 2 * nq, nu, oracle_params(d, dt), x0)
    for ob in (a, b):
        ob.set_candidate(xs, us, False)
        ob.calc()
        ob.calcdiff()
    for t in range(T + 1):
        for name in ("node_cost", "xnext", "Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu"):
            for prob in range(B):
                _close(a.get(name, prob, t), b.get(name, prob, t), "%s t=%d" % (name, t), rtol=1e-12)


@pytest.mark.parametrize("kind", ["ddp", "fddp"])
@pytest.mark.parametrize("drift_free", [True, False])
def test_solve_of_the_mapping_equals_the_two_step_model(kind, drift_free):
    nq, nu, B, T, dt = 4, 2, 3, 15, 0.02
    d, x0 = random_difflqr(nq, nu, B, 9, drift_free)
    run, term = cb.IntegratedActionModelEuler(d, dt), cb.IntegratedActionModelEuler(d, 0.)
    a = ol.OracleBatch("lqr", kind, T, 2 * nq, nu, run.params(), x0, tparams=term.params())
    b = ol.OracleBatch("euler_difflqr", kind, T, 2 * nq, nu, oracle_params(d, dt), x0, tparams=oracle_params(d, 0.))
    xs = np.repeat(x0[:, None, :], T + 1, 1)
    us = np.zeros((B, T, nu))
    for ob in (a, b):
        ob.solve(xs, us, 20, False, 1e-9)
    sa, sb = a.scalars(), b.scalars()
    np.testing.assert_array_equal(sa["n_records"], sb["n_records"])
    np.testing.assert_array_equal(sa["converged"], sb["converged"])
    _close(sa["cost"], sb["cost"], "cost")
    for name in ("xs", "us", "K", "k", "Vx", "Vxx"):
        for prob in range(B):
            _close(a.get_all(name, prob), b.get_all(name, prob), name)


def test_integrated_euler_of_free_forward_dynamics_is_the_arm_model():
    """IntegratedActionModelEuler(DifferentialActionModelFreeFwdDynamics(...), dt) — the composition the reference's arm
    factory writes (benchmark/factory/arm.hpp:76-90) — gives the arm node model with the differential model's options."""
    chain = cb.synthetic_arm(7)
    arm = np.linspace(0.1, 0.7, 7)
    dam = cb.DifferentialActionModelFreeFwdDynamics(chain, armature=arm, goal="translation", w_xReg=1e-3)
    run = cb.IntegratedActionModelEuler(dam, 2e-3)
    term = cb.IntegratedActionModelEuler(dam, 0.)
    assert isinstance(run, cb.ArmActionModel) and run.differential is dam
    assert run.dt == 2e-3 and term.dt == 0. and run.nx == 14 and run.nu == 7
    p = run.params()
    assert p.size == 36 + 28 * 7 and p[0] == 2e-3 and p[32] == 1. and p[5] == 1e-3
    np.testing.assert_array_equal(p[36 + 25::28], arm)
    np.testing.assert_array_equal(term.params()[1:], p[1:])  # the same differential model, only dt differs
    # the oracle accepts the block (layout agreed between the product's host side and the checker)
    x0 = np.zeros((1, 14))
    ob = ol.OracleBatch("arm", "fddp", 2, 14, 7, p, x0, tparams=term.params())
    ob.set_candidate(np.zeros((1, 3, 14)), np.zeros((1, 2, 7)), False)
    assert np.isfinite(ob.calc()).all()
