"""ctypes binding of the CPU oracle (oracle/_build/liboracle.so). TEST INFRASTRUCTURE ONLY.

The oracle restates the reference's SolverDDP/SolverFDDP path (see oracle/*.hpp headers for the
reference file:line each function follows). Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference leg may import this module.
"""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KIND = {"lqr": 0, "unicycle": 1, "arm": 2, "euler_difflqr": 3}
SOLVER = {"ddp": 0, "fddp": 1, "boxddp": 2, "boxfddp": 3}
SCALARS = ["cost", "stop", "d0", "d1", "xreg", "ureg", "steplength", "dV", "dVexp", "iter", "is_feasible",
           "converged", "n_records", "cost_try", "was_feasible", "_"]
TRACE = ["iter", "cost", "stop", "d0", "d1", "xreg", "ureg", "steplength", "is_feasible", "dV", "dVexp",
         "alpha_idx", "retries"]

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)


def _p(a):
    return None if a is None else a.ctypes.data_as(_dp)


def load(fast=False):
    path = os.path.join(ROOT, "oracle", "_build", "liboracle_fast.so" if fast else "liboracle.so")
    lib = C.CDLL(path)
    lib.orc_batch_create.restype = C.c_void_p
    lib.orc_batch_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _dp, C.c_long, C.c_int, _dp]
    lib.orc_batch_create2.restype = C.c_void_p
    lib.orc_batch_create2.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _dp, C.c_long, C.c_int, _dp, _dp]
    lib.orc_batch_destroy.argtypes = [C.c_void_p]
    lib.orc_batch_set.argtypes = [C.c_void_p, C.c_char_p, C.c_double]
    lib.orc_batch_set_alphas.argtypes = [C.c_void_p, _dp, C.c_int]
    lib.orc_batch_solve.argtypes = [C.c_void_p, _dp, _dp, C.c_int, C.c_int, C.c_double, C.c_int, _dp]
    lib.orc_batch_set_candidate.argtypes = [C.c_void_p, _dp, _dp, C.c_int]
    lib.orc_batch_calc.argtypes = [C.c_void_p, _dp]
    lib.orc_batch_calcdiff.argtypes = [C.c_void_p, _dp]
    lib.orc_batch_backward.argtypes = [C.c_void_p, _ip]
    lib.orc_batch_try_step.argtypes = [C.c_void_p, C.c_double, _dp, _ip]
    lib.orc_batch_expected_improvement.argtypes = [C.c_void_p, _dp, _dp]
    lib.orc_get.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_int, _dp]
    lib.orc_batch_scalars.argtypes = [C.c_void_p, _dp]
    lib.orc_get_trace.argtypes = [C.c_void_p, C.c_int, _dp, C.c_int]
    lib.orc_kkt_solve.argtypes = [C.c_void_p, C.c_int, _dp, _dp, C.c_int, _dp, _dp]
    lib.orc_numdiff.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, _dp, _dp, _dp]
    lib.orc_arm_op.argtypes = [_dp, C.c_long, C.c_char_p, _dp, _dp]
    lib.orc_batch_set_u_bounds.argtypes = [C.c_void_p, _dp, _dp, C.c_int]
    lib.orc_boxqp_solve.argtypes = [C.c_int, _dp, _dp, _dp, _dp, _dp, C.c_int, C.c_double, C.c_double, C.c_double, _dp,
                                    _dp, _ip, _ip]
    return lib


class OracleBatch:
    """A batch of B independent OCPs solved by the CPU oracle (one solver object per problem)."""

    def __init__(self, kind, solver, T, nx, nu, params, x0s, fast=False, tparams=None):
        self.lib = load(fast)
        x0s = np.ascontiguousarray(x0s, dtype=np.float64).reshape(-1, nx)
        params = np.ascontiguousarray(params, dtype=np.float64)
        self.B, self.T, self.nx, self.nu = x0s.shape[0], T, nx, nu
        per = params.ndim == 2
        nparams = params.shape[-1]
        if tparams is not None:
            tparams = np.ascontiguousarray(tparams, dtype=np.float64)
            assert tparams.shape == params.shape, "terminal block: same layout and per-problem-ness as the running one"
        self.h = self.lib.orc_batch_create2(KIND[kind], SOLVER[solver], self.B, T, nx, nu, _p(params), nparams,
                                            int(per), _p(x0s), _p(tparams))
        if not self.h:
            raise RuntimeError("orc_batch_create failed")
        self._keep = (params, x0s, tparams)

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.orc_batch_destroy(self.h)
            self.h = None

    def set(self, name, v):
        assert self.lib.orc_batch_set(self.h, name.encode(), float(v)) == 0

    def set_alphas(self, alphas):
        a = np.ascontiguousarray(alphas, dtype=np.float64)
        assert self.lib.orc_batch_set_alphas(self.h, _p(a), a.size) == 0

    def set_u_bounds(self, lb, ub):
        """ActionModelAbstract::set_u_lb / set_u_ub on the running model: [nu] shared, or [B, nu] (per-problem models)."""
        lb = np.ascontiguousarray(lb, dtype=np.float64)
        ub = np.ascontiguousarray(ub, dtype=np.float64)
        assert self.lib.orc_batch_set_u_bounds(self.h, _p(lb), _p(ub), int(lb.ndim == 2)) == 0

    def solve(self, init_xs=None, init_us=None, maxiter=100, is_feasible=False, reginit=1e-9, nthreads=0):
        xs = None if init_xs is None else np.ascontiguousarray(init_xs, dtype=np.float64)
        us = None if init_us is None else np.ascontiguousarray(init_us, dtype=np.float64)
        sec = C.c_double(0)
        n = self.lib.orc_batch_solve(self.h, _p(xs), _p(us), maxiter, int(is_feasible), reginit, nthreads,
                                     C.byref(sec))
        self.seconds = sec.value
        return n

    def set_candidate(self, xs, us, is_feasible=False):
        xs = None if xs is None else np.ascontiguousarray(xs, dtype=np.float64)
        us = None if us is None else np.ascontiguousarray(us, dtype=np.float64)
        self.lib.orc_batch_set_candidate(self.h, _p(xs), _p(us), int(is_feasible))

    def calc(self):
        c = np.zeros(self.B)
        self.lib.orc_batch_calc(self.h, _p(c))
        return c

    def calcdiff(self):
        c = np.zeros(self.B)
        self.lib.orc_batch_calcdiff(self.h, _p(c))
        return c

    def backward(self):
        st = np.zeros(self.B, dtype=np.int32)
        self.lib.orc_batch_backward(self.h, st.ctypes.data_as(_ip))
        return st

    def try_step(self, alpha):
        dV = np.zeros(self.B)
        st = np.zeros(self.B, dtype=np.int32)
        self.lib.orc_batch_try_step(self.h, alpha, _p(dV), st.ctypes.data_as(_ip))
        return dV, st

    def expected_improvement(self):
        d = np.zeros((self.B, 2))
        stop = np.zeros(self.B)
        self.lib.orc_batch_expected_improvement(self.h, _p(d), _p(stop))
        return d, stop

    _SHAPES = None

    def shape(self, name):
        n, m = self.nx, self.nu
        return {"xs": (n,), "xs_try": (n,), "fs": (n,), "Vx": (n,), "Vxx": (n, n), "Fx": (n, n), "Fu": (n, m),
                "Lx": (n,), "Lu": (m,), "Lxx": (n, n), "Lxu": (n, m), "Luu": (m, m), "xnext": (n,),
                "node_cost": (1,), "us": (m,), "us_try": (m,), "K": (m, n), "k": (m,), "Qxx": (n, n),
                "Qxu": (n, m), "Quu": (m, m), "Qx": (n,), "Qu": (m,), "Quuk": (m,)}[name]

    def get(self, name, b, t):
        shp = self.shape(name)
        out = np.zeros(int(np.prod(shp)))
        n = self.lib.orc_get(self.h, name.encode(), b, t, _p(out))
        if n < 0:
            raise KeyError((name, b, t))
        return out.reshape(shp, order="F")

    def get_all(self, name, b=None):
        """Stack field over t (and over b if b is None): shape [B,] nt, *shape."""
        terminal_too = name in ("xs", "xs_try", "fs", "Vx", "Vxx", "Lx", "Lxx", "node_cost", "Fx")
        nt = self.T + 1 if terminal_too else self.T
        if b is None:
            return np.stack([self.get_all(name, bb) for bb in range(self.B)])
        return np.stack([self.get(name, b, t) for t in range(nt)])

    def scalars(self):
        out = np.zeros((self.B, 16))
        self.lib.orc_batch_scalars(self.h, _p(out))
        return {k: out[:, i].copy() for i, k in enumerate(SCALARS) if k != "_"}

    def trace(self, b):
        n = self.lib.orc_get_trace(self.h, b, None, 0)
        out = np.zeros((max(n, 1), 13))
        self.lib.orc_get_trace(self.h, b, _p(out), n)
        return out[:n]

    def kkt_solve(self, b, init_xs, init_us, maxiter=100):
        xs = np.ascontiguousarray(init_xs, dtype=np.float64)
        us = np.ascontiguousarray(init_us, dtype=np.float64)
        xo = np.zeros((self.T + 1, self.nx))
        uo = np.zeros((self.T, self.nu))
        ok = self.lib.orc_kkt_solve(self.h, b, _p(xs), _p(us), maxiter, _p(xo), _p(uo))
        return ok, xo, uo

    def numdiff(self, b, x, u, terminal=False):
        n, m = self.nx, self.nu
        cnt = n * n + n * m + n + m
        a = np.zeros(cnt)
        d = np.zeros(cnt)
        x = np.ascontiguousarray(x, dtype=np.float64)
        u = np.ascontiguousarray(u, dtype=np.float64)
        self.lib.orc_numdiff(self.h, b, int(terminal), _p(x), _p(u), _p(a), _p(d))

        def split(v):
            return {"Fx": v[:n * n].reshape((n, n), order="F"), "Fu": v[n * n:n * n + n * m].reshape((n, m), order="F"),
                    "Lx": v[n * n + n * m:n * n + n * m + n], "Lu": v[n * n + n * m + n:]}

        return split(a), split(d)


def boxqp_solve(H, q, lb, ub, xinit, maxiter=100, th_acceptstep=0.1, th_grad=1e-9, reg=1e-9):
    """BoxQP(nx, maxiter, th_acceptstep, th_grad, reg).solve(H, q, lb, ub, xinit) of the oracle (box-qp.cpp:52-186).
    -> dict(x, Hff_inv, clamped (bool mask), iters); raises ArithmeticError on "backward_error"."""
    lib = load(False)
    n = len(q)
    f = lambda a: np.ascontiguousarray(a, dtype=np.float64)
    Hc = np.ascontiguousarray(np.asarray(H, dtype=np.float64).T)  # column-major
    x, Hi = np.zeros(n), np.zeros(n * n)
    nf, cl = C.c_int(0), np.zeros(n, dtype=np.int32)
    it = lib.orc_boxqp_solve(n, _p(Hc), _p(f(q)), _p(f(lb)), _p(f(ub)), _p(f(xinit)), maxiter, th_acceptstep, th_grad,
                             reg, _p(x), _p(Hi), C.byref(nf), cl.ctypes.data_as(_ip))
    if it < 0:
        raise ArithmeticError("backward_error")
    k = nf.value
    return dict(x=x, Hff_inv=Hi[:k * k].reshape(k, k, order="F"), clamped=cl.astype(bool), iters=it)


def lqr_params(n, m, drift_free=True, Fx=None, Fu=None, f0=None, Lxx=None, Lxu=None, Luu=None, lx=None, lu=None):
    """Pack ActionModelLQR parameters (defaults = reference ctor, lqr.hxx:14-25)."""
    Fx = np.eye(n) if Fx is None else Fx
    Fu = np.eye(n, m) if Fu is None else Fu
    f0 = np.ones(n) if f0 is None else f0
    Lxx = np.eye(n) if Lxx is None else Lxx
    Lxu = np.eye(n, m) if Lxu is None else Lxu
    Luu = np.eye(m) if Luu is None else Luu
    lx = np.ones(n) if lx is None else lx
    lu = np.ones(m) if lu is None else lu
    f = lambda a: np.asarray(a, dtype=np.float64).ravel(order="F")
    return np.concatenate([[1.0 if drift_free else 0.0], f(Fx), f(Fu), f(f0), f(Lxx), f(Lxu), f(Luu), f(lx), f(lu)])


def unicycle_params(dt=0.1, wx=10.0, wu=1.0):
    return np.array([dt, wx, wu], dtype=np.float64)


def arm_op(params, op, inp, nout):
    """Rigid-body primitive of the arm oracle by name (oracle/arm.hpp arm_op)."""
    lib = load(False)
    params = np.ascontiguousarray(params, dtype=np.float64)
    inp = np.ascontiguousarray(inp, dtype=np.float64)
    out = np.zeros(nout)
    n = lib.orc_arm_op(_p(params), params.size, op.encode(), _p(inp), _p(out))
    if n < 0:
        raise RuntimeError("orc_arm_op(%s) failed" % op)
    return out[:n]
