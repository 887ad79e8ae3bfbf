"""Development experiment: kernel time of the unicycle FDDP batch as a function of maxiter (time profile of the solve)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import crocoddyl_b200 as cb
B, T = int(sys.argv[1]) if len(sys.argv) > 1 else 65536, 100
rng = np.random.default_rng(0)
x0 = np.stack([rng.uniform(-2, 2, B), rng.uniform(-2, 2, B), rng.uniform(-np.pi, np.pi, B)], 1)
xs = np.repeat(x0[:, None, :], T + 1, 1); us = np.zeros((B, T, 2))
m = cb.ActionModelUnicycle(); p = cb.ShootingProblem(x0, [m] * T, m); s = cb.SolverFDDP(p); s.setStorage(False)
for mi in (100, 5, 10, 20, 30, 40, 60, 80, 100):
    s.solve(xs, us, mi, False, 1e-9)
    sg = s.scalars()
    print("maxiter=%3d kernel_ms=%7.2f still_running=%6d iters_done_total=%d" % (mi, s.timing()["kernel_ms"], int(np.sum((sg["converged"] == 0) & (sg["n_records"] >= mi))), int(sg["n_records"].sum())), flush=True)
print(np.percentile(sg["n_records"], [10, 25, 50, 75, 90, 95, 97, 99]))
