"""The parity build of the library (crocoddyl_b200.build.build_parity: the same sources with -fmad=false and IEEE
sqrt / divisions by the LLT pivots, like the oracle's -ffp-contract=off) against the CPU oracle, BIT for BIT.

What it establishes: on the paths whose summation order is the reference's (LQR through the warp-per-problem backward
pass, n <= 16) the CUDA solve reproduces the oracle exactly — trajectories, gains, value function, iteration counts,
step lengths, regularisation — so every difference of the DEFAULT build on those paths is FMA contraction / reciprocal pivots and
nothing else. n = 36 goes through the FP64 tensor-core tiles, which sum in another order: equal to ~1e-13, not to the
bit. The unicycle adds CUDA's sin/cos, which differ from glibc's in the last bit for most arguments: a minority of the
problems are bit-identical, all have the oracle's iteration count on this short horizon."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PARITY_LIB = os.path.join(ROOT, "crocoddyl_b200", "lib", "libcrocoddyl_b200_parity.so")


@pytest.mark.gpu
def test_parity_build_is_bit_identical_to_the_oracle_on_lqr():
    if not os.path.exists(PARITY_LIB):
        pytest.skip("parity build not present (python -m crocoddyl_b200.build --parity)")
    env = dict(os.environ, CROCODDYL_B200_LIB=PARITY_LIB)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "parity_build_check.py")], env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    res = json.loads(r.stdout.strip().splitlines()[-1])
    assert res["lib"] == "libcrocoddyl_b200_parity.so"
    for n, m in ((6, 2), (12, 4), (16, 8)):
        for kind in ("ddp", "fddp"):
            case = res["lqr_%d_%d_%s" % (n, m, kind)]
            for f in ("xs", "us", "K", "k", "Vx", "Vxx", "decisions"):
                assert case[f], (n, m, kind, f, case["max_abs_diff"])
            assert case["cost_rel_diff"] < 1e-14  # the total cost is re-summed by the scalar export, not bit-identical
    assert res["lqr_36_12_fddp"]["max_abs_diff"] < 1e-11  # tensor-core tiles: another summation order
    uni = res["unicycle"]
    assert uni["same_iterations"] == uni["B"]
    assert uni["max_abs_diff"] < 1e-12
    assert uni["bitwise_equal_problems"] >= 1
