"""Shared by the CPU and GPU tests of IntegratedActionModelEuler(DifferentialActionModelLQR): a random differential model
(well-conditioned cost, per-problem matrices) and its parameter block for the oracle's two-step statement
(oracle/models.hpp ActionModelEulerDiffLQR: diff-lqr.hxx then euler.hxx)."""
import numpy as np

import crocoddyl_b200 as cb


def random_difflqr(nq, nu, B, seed, drift_free):
    rng = np.random.default_rng(seed)
    nx = 2 * nq
    d = cb.DifferentialActionModelLQR(nq, nu, drift_free)
    d.Fq = -2.0 * np.eye(nq) + 0.3 * rng.standard_normal((B, nq, nq))
    d.Fv = -0.5 * np.eye(nq) + 0.3 * rng.standard_normal((B, nq, nq))
    d.Fu = rng.standard_normal((B, nq, nu))
    d.f0 = rng.standard_normal((B, nq))
    A = rng.standard_normal((B, nx + nu, nx + nu))
    L = np.einsum("bki,bkj->bij", A, A) / (nx + nu) + np.eye(nx + nu)
    d.Lxx, d.Lxu, d.Luu = L[:, :nx, :nx], L[:, :nx, nx:], L[:, nx:, nx:]
    d.lx = rng.standard_normal((B, nx))
    d.lu = rng.standard_normal((B, nu))
    x0 = rng.standard_normal((B, nx))
    return d, x0


def oracle_params(d, dt):
    """[dt, drift_free, Fq, Fv, Fu, f0, Lxx, Lxu, Luu, lx, lu] per problem, matrices column-major."""
    B = np.asarray(d.Fq).shape[0]
    cols = [np.full((B, 1), dt), np.full((B, 1), 1.0 if d.driftFree else 0.0)]
    for name in ("Fq", "Fv", "Fu", "f0", "Lxx", "Lxu", "Luu", "lx", "lu"):
        a = np.asarray(getattr(d, name), dtype=np.float64)
        cols.append(np.transpose(a, (0,) + tuple(range(a.ndim - 1, 0, -1))).reshape(B, -1))
    return np.ascontiguousarray(np.concatenate(cols, 1))
