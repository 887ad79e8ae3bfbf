"""Development timing: phases of one FDDP iteration of the arm batch through the step-wise API."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import crocoddyl_b200 as cb
T = 100
model = cb.ArmActionModel(cb.synthetic_arm(7), 1e-3, w_gripperPose=1., w_xReg=1e-4, w_uReg=1e-4)
for B in [int(a) for a in (sys.argv[1:] or ["1", "148", "1024"])]:
    rng = np.random.default_rng(2)
    x0 = np.concatenate([rng.uniform(-1, 1, (B, 7)), np.zeros((B, 7))], 1)
    xs = np.repeat(x0[:, None, :], T + 1, 1)
    problem = cb.ShootingProblem(x0, [model] * T, model.terminal_model)
    us = np.repeat(problem.quasiStatic(x0)[:, None, :], T, 1)
    solver = cb.SolverFDDP(problem)
    solver.setStorage(False)
    solver.setCandidate(xs, us, False)
    solver.x_reg = solver.u_reg = 0.1
    out = {}
    for rep in range(2):
        solver.computeDirection(True); out["calc+calcdiff+backward"] = solver.timing()["kernel_ms"]
        solver.computeDirection(False); out["backward"] = solver.timing()["kernel_ms"]
        solver.tryStep(1.0); out["forward"] = solver.timing()["kernel_ms"]
    print("B=%d " % B + " ".join("%s=%.3fms" % kv for kv in out.items()), flush=True)
