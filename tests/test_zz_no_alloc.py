"""cgx_unicycle_calc_noalloc (no counterpart in the reference): the
allocation-free variant of ShootingProblem::calc for unicycle batches, whose scratch is a static device array. Compared,
through the C ABI, with the existing entry point (cgx_problem_calc behind ShootingProblem.calc) on two calls in a row —
the second with other inputs and another size, because the array keeps the first call's contents. Collected last
(test_zz_*). Passes alone in a fresh process (the entry point needs no handle) and
after the other tests."""
import ctypes as C

import numpy as np
import pytest

import crocoddyl_b200 as cb
from crocoddyl_b200 import _lib

_dp = C.POINTER(C.c_double)


def _p(a):
    return None if a is None else a.ctypes.data_as(_dp)


def _inputs(B, T, seed):
    rng = np.random.default_rng(seed)
    xs = np.concatenate([rng.uniform(-2, 2, (B, T + 1, 2)), rng.uniform(-np.pi, np.pi, (B, T + 1, 1))], 2)
    us = rng.uniform(-1, 1, (B, T, 2))
    return np.ascontiguousarray(xs), np.ascontiguousarray(us)


def _noalloc(B, T, params, tparams, xs, us):
    costs = np.full(B, np.nan)
    rc = cb.lib().cgx_unicycle_calc_noalloc(B, T, _p(params), _p(tparams), _p(xs), _p(us), _p(costs))
    return rc, costs


def test_sizes_beyond_the_static_scratch_are_rejected():
    """Argument checks come before any device work, so this half also runs on the CPU-only box."""
    T = 100
    B = _lib.NOALLOC_SCRATCH_DOUBLES // (6 * T + 5) + 1
    params = np.array([0.1, 10., 1.])
    xs, us = np.zeros((B, T + 1, 3)), np.zeros((B, T, 2))
    rc, _ = _noalloc(B, T, params, None, xs, us)
    assert rc == -1
    msg = cb.lib().cgx_last_error().decode()
    assert "CGX_NOALLOC_SCRATCH_DOUBLES" in msg and str(B * (6 * T + 5)) in msg
    rc, _ = _noalloc(0, T, params, None, xs, us)
    assert rc == -1
    assert cb.lib().cgx_unicycle_calc_noalloc(4, T, None, None, _p(xs), _p(us), _p(np.zeros(4))) == -1


@pytest.mark.gpu
def test_noalloc_calc_equals_problem_calc_on_two_calls_in_a_row():
    run = cb.ActionModelUnicycle()
    term = cb.ActionModelUnicycle()
    term.costWeights = np.array([30., 2.])  # a terminal model of its own: exercises tparams
    # call 1: the largest batch that fits at T = 100; call 2: other inputs, smaller and ragged, so stale words of call 1
    # sit right behind (and inside the old layout of) everything call 2 uses
    for B, T, seed, terminal in ((1733, 100, 1, run), (333, 37, 2, term)):
        xs, us = _inputs(B, T, seed)
        problem = cb.ShootingProblem(xs[:, 0].copy(), [run] * T, terminal)
        want = problem.calc(xs, us)  # existing entry point: cgx_problem_calc
        rc, got = _noalloc(B, T, run.params(), None if terminal is run else terminal.params(), xs, us)
        assert rc == 0, cb.lib().cgx_last_error().decode()
        assert np.all(np.isfinite(got))
        np.testing.assert_allclose(got, want, rtol=1e-9, atol=1e-9)  # the tolerance of the parity tests (same arithmetic)
        del problem
    # and the first call again, after the second one overwrote the array: same answer as before
    xs, us = _inputs(1733, 100, 1)
    rc, again = _noalloc(1733, 100, run.params(), None, xs, us)
    assert rc == 0
    problem = cb.ShootingProblem(xs[:, 0].copy(), [run] * 100, run)
    np.testing.assert_allclose(again, problem.calc(xs, us), rtol=1e-9, atol=1e-9)
