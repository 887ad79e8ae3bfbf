"""The CUDA path (through the C ABI) against the committed golden vectors; needs neither the oracle nor /root/reference.
Discrete decisions (iterations, accepted step length index, regularisation, feasibility per iteration) must be
identical; cost, xs, us, K, k, Vx, Vxx within 1e-9 relative (BASELINE.json north_star), 1e-7 on the trajectories of
runs with tens of iterations where per-iteration 1e-12 differences compound."""
import os

import numpy as np
import pytest

import crocoddyl_b200 as cb
from golden import cases

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _close(a, b, name, rtol):
    scale = max(1.0, float(np.max(np.abs(b))))
    np.testing.assert_allclose(a, b, rtol=rtol, atol=rtol * scale, err_msg=name)


@pytest.mark.parametrize("name", cases.CASES)
def test_solve_matches_golden(name):
    gold = np.load(os.path.join(GOLD, name + ".npz"))
    cs = cases.build(name, cb)
    T = cs["T"]
    problem = cb.ShootingProblem(cs["x0"], [cs["model"]] * T, cs["terminal"])
    us = cs["us"]
    if us is None:
        us = np.repeat(problem.quasiStatic(cs["x0"])[:, None, :], T, 1)
        _close(us, gold["init_us"], "quasiStatic", 1e-9)
    solver = (cb.SolverFDDP if cs["solver"] == "fddp" else cb.SolverDDP)(problem)
    solver.enableTrace(cs["maxiter"] + 1)
    solver.solve(cs["xs"], us, cs["maxiter"], False, cs["reginit"])
    sg = solver.scalars()
    assert np.array_equal(sg["iter"], gold["s_iter"]), (sg["iter"], gold["s_iter"])
    assert np.array_equal(sg["converged"], gold["s_converged"])
    assert np.array_equal(sg["n_records"], gold["s_n_records"])
    _close(sg["cost"], gold["s_cost"], "cost", 1e-9)
    rtol = 1e-9 if int(np.max(gold["s_iter"])) <= 2 else 1e-7
    for f in cases.FIELDS:
        _close(solver.get(f), gold[f], f, rtol)
    for b in range(cs["x0"].shape[0]):
        tg, to = solver.trace(b), gold["trace_%d" % b]
        assert tg.shape == to.shape
        for col in cases.TRACE_EXACT:
            assert np.array_equal(tg[:, col], to[:, col]), (b, col)
        np.testing.assert_allclose(tg[:, 1], to[:, 1], rtol=1e-9)
