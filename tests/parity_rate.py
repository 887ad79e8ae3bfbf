"""Test-infrastructure script (not collected by pytest): iteration-count / accept-sequence agreement rate of the
CUDA unicycle FDDP path with the CPU oracle on the first N problems of config C2."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import crocoddyl_b200 as cb
import oracle_lib as ol

N = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
T = 100
rng = np.random.default_rng(0)
x0 = np.stack([rng.uniform(-2, 2, 65536), rng.uniform(-2, 2, 65536), rng.uniform(-np.pi, np.pi, 65536)], 1)[:N]
xs = np.repeat(x0[:, None, :], T + 1, 1)
us = np.zeros((N, T, 2))
m = cb.ActionModelUnicycle()
p = cb.ShootingProblem(x0, [m] * T, m)
s = cb.SolverFDDP(p)
s.setStorage(False)
conv = s.solve(xs, us, 100, False, 1e-9)
sg = s.scalars()
ob = ol.OracleBatch("unicycle", "fddp", T, 3, 2, ol.unicycle_params(), x0, fast=False)
ob.solve(xs, us, 100, False, 1e-9)
so = ob.scalars()
same = so["n_records"] == sg["n_records"]
xg, xo = s.get("xs"), ob.get_all("xs")
err = np.abs(xg - xo).reshape(N, -1).max(1)
print("lib=%s N=%d same_iters=%.5f (%d differ) | of the oracle-converged in <100 iters: %.5f | max|dxs| where same: %.3e | kernel_ms=%.2f"
      % (os.path.basename(cb._lib.LIB_PATH), N, same.mean(), (~same).sum(),
         same[(so["converged"] == 1) & (so["n_records"] < 100)].mean(), err[same & (so["converged"] == 1)].max(),
         s.timing()["kernel_ms"]))
d = np.where(~same)[0][:10]
print(" differing problems (oracle iters, gpu iters):", [(int(i), int(so["n_records"][i]), int(sg["n_records"][i])) for i in d])
