"""Development experiment: forward/backward sweep counts of the unicycle OCPs by iteration count."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import crocoddyl_b200 as cb
B, T = 65536, 100
rng = np.random.default_rng(0)
x0 = np.stack([rng.uniform(-2, 2, B), rng.uniform(-2, 2, B), rng.uniform(-np.pi, np.pi, B)], 1)
xs = np.repeat(x0[:, None, :], T + 1, 1); us = np.zeros((B, T, 2))
m = cb.ActionModelUnicycle(); p = cb.ShootingProblem(x0, [m] * T, m); s = cb.SolverFDDP(p); s.setStorage(False)
s.solve(xs, us, 100, False, 1e-9)
sg = s.scalars()
it, nb, nf = sg["n_records"], sg["n_backward"], sg["n_forward"]
for lo, hi in ((0, 20), (20, 40), (40, 60), (60, 80), (80, 100), (100, 101)):
    k = (it >= lo) & (it < hi)
    if k.sum():
        print("iters [%3d,%3d): %6d problems  fwd/iter=%.2f bwd/iter=%.2f  sum fwd=%d" % (lo, hi, k.sum(), (nf[k] / it[k]).mean(), (nb[k] / it[k]).mean(), nf[k].sum()))
