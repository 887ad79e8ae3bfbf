"""Development timing of the arm FDDP solve kernel (not a bench line): prints kernel ms for a few batch sizes."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import crocoddyl_b200 as cb  # noqa: E402

T = 100
model = cb.ArmActionModel(cb.synthetic_arm(7), 1e-3, w_gripperPose=1., w_xReg=1e-4, w_uReg=1e-4)
for B in [int(a) for a in (sys.argv[1:] or ["1024", "8192"])]:
    rng = np.random.default_rng(2)
    x0 = np.concatenate([rng.uniform(-1, 1, (B, 7)), np.zeros((B, 7))], 1)
    xs = np.repeat(x0[:, None, :], T + 1, 1)
    problem = cb.ShootingProblem(x0, [model] * T, model.terminal_model)
    us = np.repeat(problem.quasiStatic(x0)[:, None, :], T, 1)
    solver = cb.SolverFDDP(problem)
    solver.setStorage(False)
    ms = []
    for _ in range(3):
        conv = solver.solve(xs, us, 100, False, 0.1)
        ms.append(solver.timing()["kernel_ms"])
    sg = solver.scalars()
    print("lib=%s B=%d kernel_ms=%s solves/s=%.0f iters mean=%.2f max=%d conv=%.3f cost_sum=%.12g" %
          (os.path.basename(cb._lib.LIB_PATH), B, ["%.1f" % m for m in ms], B / (min(ms) * 1e-3), sg["n_records"].mean(),
           sg["n_records"].max(), conv.mean(), sg["cost"].sum()), flush=True)
