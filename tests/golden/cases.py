"""Golden cases of the solve path: small seeded problems whose oracle outputs are committed as tests/golden/*.npz.

The reference holds no numeric golden vectors for this path and cannot be built or imported in this image,
so these vectors are generated from the CPU oracle (tests/golden/make_golden.py) after the oracle has been pinned by the
reference's own known-answer structures (tests/test_oracle_cpu.py, tests/test_oracle_arm_cpu.py). They freeze the
oracle (CPU test) and give the GPU box a fixture that needs neither the oracle build nor /root/reference.
Shapes follow the reference's tests: unittest/test_solvers.cpp:65-110 (LQR +- drift, T = 10), unicycle T = 10 from
xs = x0, benchmark/factory/arm.hpp:60-90 (arm, quasi-static warm start, reginit 0.1).
"""
import numpy as np

FIELDS = ("xs", "us", "K", "k", "Vx", "Vxx")
TRACE_EXACT = (0, 5, 6, 7, 8, 11, 12)  # iter, xreg, ureg, steplength, is_feasible, accepted alpha, retries


def _lqr(cb, n, m, B, seed, drift_free):
    rng = np.random.default_rng(seed)
    mod = cb.ActionModelLQR(n, m, drift_free)
    mod.Fx = np.eye(n) + 0.1 * rng.standard_normal((B, n, n)) / np.sqrt(n)
    mod.Fu = rng.standard_normal((B, n, m))
    A = rng.standard_normal((B, n, n))
    Bm = rng.standard_normal((B, m, m))
    mod.Lxx = np.einsum("bki,bkj->bij", A, A) / n + np.eye(n)
    mod.Luu = np.einsum("bki,bkj->bij", Bm, Bm) / m + np.eye(m)
    mod.Lxu = 0.1 * rng.standard_normal((B, n, m))
    mod.lx = rng.standard_normal((B, n))
    mod.lu = rng.standard_normal((B, m))
    mod.f0 = rng.standard_normal((B, n))
    return mod, rng.standard_normal((B, n))


def build(name, cb):
    """-> dict(kind, solver, model, terminal, x0, xs, us, T, reginit, maxiter); us None = quasiStatic(x0)."""
    if name.startswith("lqr"):
        drift_free = name.endswith("driftfree")
        B, T, n, m = 3, 10, 8, 4
        mod, x0 = _lqr(cb, n, m, B, 11, drift_free)
        return dict(kind="lqr", solver=name.split("_")[1], model=mod,
                    terminal=mod, x0=x0, xs=np.repeat(x0[:, None, :], T + 1, 1), us=np.zeros((B, T, m)), T=T,
                    reginit=1e-9, maxiter=10)
    if name.startswith("unicycle"):
        B, T = 8, 20
        rng = np.random.default_rng(5)
        x0 = np.stack([rng.uniform(-2, 2, B), rng.uniform(-2, 2, B), rng.uniform(-np.pi, np.pi, B)], 1)
        mod = cb.ActionModelUnicycle()
        return dict(kind="unicycle", solver=name.split("_")[1], model=mod, terminal=mod, x0=x0,
                    xs=np.repeat(x0[:, None, :], T + 1, 1), us=np.zeros((B, T, 2)), T=T, reginit=1e-9, maxiter=100)
    if name.startswith("arm"):
        B, T, nv = 3, 12, 7
        mod = cb.ArmActionModel(cb.synthetic_arm(nv), 1e-3, w_gripperPose=1., w_xReg=1e-4, w_uReg=1e-4)
        rng = np.random.default_rng(2)
        x0 = np.concatenate([rng.uniform(-1, 1, (B, nv)), np.zeros((B, nv))], 1)
        return dict(kind="arm", solver=name.split("_")[1], model=mod, terminal=mod.terminal_model,
                    x0=x0, xs=np.repeat(x0[:, None, :], T + 1, 1), us=None, T=T, reginit=0.1, maxiter=100)
    raise KeyError(name)


CASES = ("lqr_ddp_driftfree", "lqr_fddp_drift", "unicycle_fddp", "unicycle_ddp", "arm_fddp", "arm_ddp")
