"""The CPU oracle against the committed golden vectors (tests/golden/*.npz, made by tests/golden/make_golden.py)."""
import os

import numpy as np
import pytest

from golden import cases
from golden.make_golden import oracle_run

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("name", cases.CASES)
def test_oracle_reproduces_golden(name):
    gold = np.load(os.path.join(GOLD, name + ".npz"))
    out = oracle_run(name)
    assert sorted(gold.files) == sorted(out.keys())
    for k in gold.files:
        a, g = np.asarray(out[k]), gold[k]
        assert a.shape == g.shape, k
        if k.startswith("trace_"):
            for col in cases.TRACE_EXACT:
                assert np.array_equal(a[:, col], g[:, col]), (k, col)
        if k in ("s_iter", "s_converged", "s_n_records", "s_is_feasible"):
            assert np.array_equal(a, g), k
        else:
            np.testing.assert_allclose(a, g, rtol=1e-10, atol=1e-10 * max(1., float(np.max(np.abs(g)))), err_msg=k)


def test_golden_cases_are_the_reference_test_shapes():
    # LQR: DDP solves it in one Newton step (unittest/test_solvers.cpp:65-110)
    for name in ("lqr_ddp_driftfree", "lqr_fddp_drift"):
        gold = np.load(os.path.join(GOLD, name + ".npz"))
        assert np.all(gold["s_iter"] <= 2) and np.all(gold["s_converged"] == 1)
    gold = np.load(os.path.join(GOLD, "arm_fddp.npz"))
    assert np.all(gold["s_converged"] == 1)
