"""Regression tests for defects found in review, through the C ABI against the CPU oracle."""
import numpy as np
import pytest

import crocoddyl_b200 as cb
import oracle_lib as ol

pytestmark = pytest.mark.gpu


def _uni(B, T, seed):
    rng = np.random.default_rng(seed)
    x0 = np.stack([rng.uniform(-2, 2, B), rng.uniform(-2, 2, B), rng.uniform(-np.pi, np.pi, B)], 1)
    return cb.ActionModelUnicycle(), ol.unicycle_params(), x0


def _lqr(B, T, seed, n=6, m=3):
    rng = np.random.default_rng(seed)
    mod = cb.ActionModelLQR(n, m, False)
    mod.Fx = np.eye(n) + 0.1 * rng.standard_normal((n, n))
    mod.Fu = rng.standard_normal((n, m))
    mod.f0 = rng.standard_normal(n)
    return mod, mod.params(), rng.standard_normal((B, n))


@pytest.mark.parametrize("family", ["unicycle", "lqr"])
def test_set_x0_reaches_a_live_solver(family):
    """problem.x0 = new; solver.solve(...) — the receding-horizon pattern (shooting.hxx:393-400) — must solve from the
    NEW initial state on every backend (the thread-per-problem backend keeps a transposed copy of x0)."""
    B, T = 37, 15
    mod, params, x0 = (_uni if family == "unicycle" else _lqr)(B, T, 3)
    problem = cb.ShootingProblem(x0, [mod] * T, mod)
    solver = cb.SolverFDDP(problem)
    xs, us = np.repeat(x0[:, None, :], T + 1, 1), np.zeros((B, T, mod.nu))
    solver.solve(xs, us, 100, False, 1e-9)
    x0b = x0 + 0.3 * np.random.default_rng(4).standard_normal(x0.shape)
    problem.x0 = x0b
    np.testing.assert_array_equal(problem.x0, x0b)
    xsb = np.repeat(x0b[:, None, :], T + 1, 1)
    conv = solver.solve(xsb, us, 100, False, 1e-9)
    fresh = cb.SolverFDDP(cb.ShootingProblem(x0b, [mod] * T, mod))
    conv_f = fresh.solve(xsb, us, 100, False, 1e-9)
    assert np.array_equal(conv, conv_f)
    np.testing.assert_array_equal(solver.xs, fresh.xs)
    np.testing.assert_array_equal(solver.us, fresh.us)
    ob = ol.OracleBatch(family, "fddp", T, mod.nx, mod.nu, params, x0b)
    ob.solve(xsb, us, 100, False, 1e-9)
    so, sg = ob.scalars(), solver.scalars()
    # solves of many tens of iterations sit on last-bit ties of the stopping test (profiles/r02_sincos_sensitivity.json);
    # the identity with the oracle is asserted on the others
    short = so["n_records"] <= 60
    assert short.sum() >= B // 2
    assert np.array_equal(so["n_records"][short], sg["n_records"][short])
    np.testing.assert_allclose(sg["cost"][short], so["cost"][short], rtol=1e-9)
    np.testing.assert_allclose(solver.xs[:, 0], x0b, atol=1e-12)
    np.testing.assert_allclose(solver.xs[short], ob.get_all("xs")[short], rtol=1e-8, atol=1e-8)
    # the problem-level rollout sees it too
    np.testing.assert_allclose(problem.rollout(us)[:, 0], x0b, atol=0)
