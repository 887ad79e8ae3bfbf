"""GPU parity of the solver's failure / edge paths against the CPU oracle (through the C ABI):
 * "backward_error": LLT of Quu fails -> increaseRegularization + retry with recalcDiff = false
   (src/core/solvers/ddp.cpp:55-68,298-310; fddp.cpp:35-48), until it succeeds or xreg hits regmax -> solve returns false;
 * "forward_error": NaN / inf / >= 1e30 in the trial cost or state -> the step length is skipped
   (ddp.cpp:78-82, fddp.cpp:58-60, solver-base.cpp:175-181), failed line search -> regularisation grows;
 * degenerate sizes: maxiter = 0, T = 1, B = 1."""
import numpy as np
import pytest

import crocoddyl_b200 as cb
import oracle_lib as ol

pytestmark = pytest.mark.gpu
DISCRETE = (0, 5, 6, 7, 8, 11, 12)  # iter, xreg, ureg, steplength, is_feasible, accepted alpha, retries


def _lqr(n, m, B, seed, luu_shift):
    rng = np.random.default_rng(seed)
    mod = cb.ActionModelLQR(n, m, True)
    mod.Fx = np.eye(n) + 0.1 * rng.standard_normal((B, n, n)) / np.sqrt(n)
    mod.Fu = 0.3 * rng.standard_normal((B, n, m))
    A = rng.standard_normal((B, n, n))
    mod.Lxx = np.einsum("bki,bkj->bij", A, A) / n + np.eye(n)
    mod.Luu = np.repeat((luu_shift * np.eye(m))[None], B, 0)
    mod.Lxu = np.zeros((B, n, m))
    mod.lx = rng.standard_normal((B, n))
    mod.lu = rng.standard_normal((B, m))
    return mod, rng.standard_normal((B, n))


def _solve_both(kind, mod, x0, T, maxiter, reginit=1e-9):
    B = x0.shape[0]
    problem = cb.ShootingProblem(x0, [mod] * T, mod)
    solver = (cb.SolverFDDP if kind == "fddp" else cb.SolverDDP)(problem)
    ob = ol.OracleBatch("lqr", kind, T, mod.nx, mod.nu, mod.params(), x0)
    xs = np.repeat(x0[:, None, :], T + 1, 1)
    us = np.zeros((B, T, mod.nu))
    solver.enableTrace(maxiter + 1)
    conv = solver.solve(xs, us, maxiter, False, reginit)
    ob.solve(xs, us, maxiter, False, reginit)
    return solver, ob, conv


def _same_discrete(solver, ob, conv, B):
    so, sg = ob.scalars(), solver.scalars()
    assert np.array_equal(so["converged"].astype(bool), np.atleast_1d(conv))
    for k in ("n_records", "iter", "is_feasible", "xreg", "ureg", "steplength"):
        assert np.array_equal(so[k], sg[k]), (k, so[k], sg[k])
    for b in range(B):
        tg, to = solver.trace(b), ob.trace(b)
        assert tg.shape == to.shape
        for col in DISCRETE:
            assert np.array_equal(tg[:, col], to[:, col]), (b, col, tg[:, col], to[:, col])
    return so, sg


@pytest.mark.parametrize("kind", ["ddp", "fddp"])
@pytest.mark.parametrize("n,m,shift", [(6, 2, -0.05), (20, 10, -0.3)])  # warp-per-OCP and block-wide backward pass
def test_llt_failure_retries_with_larger_regularisation(kind, n, m, shift):
    # Luu = shift * I < 0: Quu = Luu + Fu' Vxx Fu is indefinite until ureg is large enough
    B, T = 4, 8
    mod, x0 = _lqr(n, m, B, 11, shift)
    solver, ob, conv = _solve_both(kind, mod, x0, T, 30)
    so, sg = _same_discrete(solver, ob, conv, B)
    assert max(solver.trace(b)[:, 12].max() for b in range(B)) > 0  # the retry path was taken
    scale = np.maximum(1., np.abs(so["cost"]))
    np.testing.assert_allclose(sg["cost"], so["cost"], rtol=1e-9, atol=1e-9 * scale.max())
    for name in ("xs", "us", "K", "k"):
        ref = ob.get_all(name)
        np.testing.assert_allclose(solver.get(name), ref, rtol=1e-8, atol=1e-8 * max(1., np.abs(ref).max()), err_msg=name)


@pytest.mark.parametrize("kind", ["ddp", "fddp"])
def test_regularisation_reaches_regmax_and_solve_returns_false(kind):
    # Luu = -1e12 I: no ureg <= regmax = 1e9 makes Quu positive definite -> solve gives up (ddp.cpp:63-65)
    B, T = 3, 5
    mod, x0 = _lqr(5, 2, B, 3, -1e12)
    solver, ob, conv = _solve_both(kind, mod, x0, T, 20)
    so, sg = _same_discrete(solver, ob, conv, B)
    assert not conv.any()
    assert np.all(sg["xreg"] == 1e9)


@pytest.mark.parametrize("kind", ["ddp", "fddp"])
def test_forward_error_skips_step_lengths(kind):
    # unicycle far from the origin: cost = 50 |x|^2 >= 1e30 -> raiseIfNaN fires for every step length
    B, T = 4, 10
    x0 = np.array([[1e16, 0., 0.], [0., -3e15, 1.], [1., 1., 0.5], [2e17, 1e17, -1.]])
    model = cb.ActionModelUnicycle()
    problem = cb.ShootingProblem(x0, [model] * T, model)
    solver = (cb.SolverFDDP if kind == "fddp" else cb.SolverDDP)(problem)
    ob = ol.OracleBatch("unicycle", kind, T, 3, 2, ol.unicycle_params(), x0)
    xs = np.repeat(x0[:, None, :], T + 1, 1)
    us = np.zeros((B, T, 2))
    solver.enableTrace(16)
    conv = solver.solve(xs, us, 12, False, 1e-9)
    ob.solve(xs, us, 12, False, 1e-9)
    _same_discrete(solver, ob, conv, B)
    assert (solver.trace(0)[:, 11] == -1).any()  # iterations without any accepted step length


def test_degenerate_sizes():
    model = cb.ActionModelUnicycle()
    x0 = np.array([[1., -1., 0.3]])
    for T, maxiter in ((1, 10), (3, 0), (2, 1)):
        problem = cb.ShootingProblem(x0, [model] * T, model)
        solver = cb.SolverFDDP(problem)
        ob = ol.OracleBatch("unicycle", "fddp", T, 3, 2, ol.unicycle_params(), x0)
        xs = np.repeat(x0[:, None, :], T + 1, 1)
        us = np.zeros((1, T, 2))
        conv = solver.solve(xs, us, maxiter, False, 1e-9)
        ob.solve(xs, us, maxiter, False, 1e-9)
        so, sg = ob.scalars(), solver.scalars()
        assert np.array_equal(so["converged"].astype(bool), np.atleast_1d(conv))
        assert np.array_equal(so["iter"], sg["iter"]) and np.array_equal(so["n_records"], sg["n_records"])
        np.testing.assert_allclose(solver.get("xs"), ob.get_all("xs"), rtol=1e-9, atol=1e-9)
        np.testing.assert_allclose(solver.get("us"), ob.get_all("us"), rtol=1e-9, atol=1e-9)
