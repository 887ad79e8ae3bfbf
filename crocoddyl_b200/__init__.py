"""crocoddyl_b200 — B200-native batched DDP/FDDP solve path behind crocoddyl's API names.

Python mirror of the reference's Python bindings for this path (bindings/python/crocoddyl/core:
actions/unicycle.cpp, actions/lqr.cpp, optctrl/shooting.cpp, solver-base.cpp, solvers/ddp.cpp,
solvers/fddp.cpp), thin over the C ABI in include/crocoddyl_b200.h. Every object is a BATCH of B
independent problems: arrays carry a leading batch dimension, B = 1 is the reference's behaviour.
All compute runs in the CUDA library; there is no CPU fallback.
"""
import ctypes as C

import numpy as np

from . import _lib
from .base import ActionDataAbstract, ActionModelAbstract, StateVector, _f64  # noqa: F401
from .arm import ArmActionModel, RevoluteChain, synthetic_arm  # noqa: F401
from ._lib import CgxError, check, dptr, lib  # noqa: F401

MODEL_LQR, MODEL_UNICYCLE, MODEL_ARM = 0, 1, 2
SOLVER_DDP, SOLVER_FDDP = 0, 1


class ActionModelUnicycle(ActionModelAbstract):
    """ActionModelUnicycle (include/crocoddyl/core/actions/unicycle.hxx:13-73): nx = 3, nu = 2, dt = 0.1,
    costWeights = (10, 1)."""
    kind = MODEL_UNICYCLE
    writes_residual = True  # data.r = [w0 x; w1 u] (unicycle.hxx:39-40)

    def __init__(self):
        self._init_abstract(3, 2, 5)
        self.dt = 0.1
        self.costWeights = np.array([10., 1.])

    def params(self):
        return np.array([self.dt, self.costWeights[0], self.costWeights[1]], dtype=np.float64)


class ActionModelLQR(ActionModelAbstract):
    """ActionModelLQR(nx, nu, driftFree=True) (include/crocoddyl/core/actions/lqr.hxx:14-70). Every matrix may
    carry a leading batch dimension ([B, ...]) to give each problem of the batch its own model."""
    kind = MODEL_LQR

    def __init__(self, nx, nu, driftFree=True):
        self._init_abstract(nx, nu, 0)
        self.driftFree = bool(driftFree)
        self.Fx = np.eye(nx)
        self.Fu = np.eye(nx, nu)
        self.f0 = np.ones(nx)
        self.Lxx = np.eye(nx)
        self.Lxu = np.eye(nx, nu)
        self.Luu = np.eye(nu)
        self.lx = np.ones(nx)
        self.lu = np.ones(nu)

    def _batch(self):
        n, m = self.nx, self.nu
        base = {"Fx": 2, "Fu": 2, "f0": 1, "Lxx": 2, "Lxu": 2, "Luu": 2, "lx": 1, "lu": 1}
        B = None
        for k, nd in base.items():
            a = np.asarray(getattr(self, k))
            if a.ndim == nd + 1:
                B = a.shape[0] if B is None else B
                if a.shape[0] != B:
                    raise ValueError("inconsistent batch dimension in ActionModelLQR." + k)
        return B

    def params(self):
        n, m = self.nx, self.nu
        B = self._batch()
        shapes = [("Fx", (n, n)), ("Fu", (n, m)), ("f0", (n,)), ("Lxx", (n, n)), ("Lxu", (n, m)), ("Luu", (m, m)),
                  ("lx", (n,)), ("lu", (m,))]
        nb = 1 if B is None else B
        cols = [np.full((nb, 1), 1.0 if self.driftFree else 0.0)]
        for k, shp in shapes:
            a = np.asarray(getattr(self, k), dtype=np.float64)
            if a.shape == shp:
                a = np.broadcast_to(a, (nb,) + shp)
            elif a.shape != (nb,) + shp:
                raise ValueError("ActionModelLQR.%s has wrong dimension (it should be %s)" % (k, shp))
            # column-major flattening of every matrix
            cols.append(np.transpose(a, (0,) + tuple(range(a.ndim - 1, 0, -1))).reshape(nb, -1))
        p = np.ascontiguousarray(np.concatenate(cols, axis=1))
        return p[0] if B is None else p


from .integrated import (DifferentialActionModelFreeFwdDynamics, DifferentialActionModelLQR,  # noqa: E402,F401
                         make_integrated_euler)

IntegratedActionModelEuler = make_integrated_euler(ActionModelLQR)


class ShootingProblem(object):
    """ShootingProblem(x0, runningModels, terminalModel) (include/crocoddyl/core/optctrl/shooting.hpp:48-63).

    x0 is [nx] (one problem) or [B, nx] (a batch). ONE running model object is aliased into all T running nodes, as
    the reference's benchmarks and examples do (benchmark/unicycle-optctrl.cpp:25-32, examples/arm_manipulation.py:63);
    the terminal model may be a different object of the same family and dimensions (its own weights, dt, ...: the
    terminal cost = gripperPose only of examples/arm_manipulation.py:41-57 is expressible). Distinct running models
    per node raise."""

    def __init__(self, x0, runningModels, terminalModel):
        if len(runningModels) == 0:
            raise ValueError("Invalid argument: at least one running model is needed")
        model = runningModels[0]
        if any(m is not model for m in runningModels):
            raise CgxError(-3, "only problems that alias one running model into every node are supported")
        if terminalModel.kind != model.kind or terminalModel.nx != model.nx or terminalModel.nu != model.nu:
            raise CgxError(-3, "the terminal model must be of the running model's family and dimensions")
        self.model = model
        self.runningModels = list(runningModels)
        self.terminalModel = terminalModel
        x0 = _f64(x0)
        if x0.ndim == 1:
            x0 = x0[None, :]
        if x0.shape[1] != model.nx:
            raise CgxError(-1, "Invalid argument: x0 has wrong dimension (it should be %d)" % model.nx)
        self.B, self.T = x0.shape[0], len(runningModels)
        self.nx, self.ndx, self.nu_max = model.nx, model.ndx, model.nu
        params = _f64(model.params())
        tparams = params if terminalModel is model else _f64(terminalModel.params())
        per = params.ndim == 2 or tparams.ndim == 2
        if per:  # per-problem blocks: both blocks get the batch dimension
            params = np.ascontiguousarray(np.broadcast_to(params, (self.B, params.shape[-1])))
            tparams = np.ascontiguousarray(np.broadcast_to(tparams, (self.B, tparams.shape[-1])))
        if per and params.shape[0] != self.B:
            raise CgxError(-1, "Invalid argument: model batch dimension differs from x0's")
        self._h = C.c_void_p()
        self._keep = (params, x0, tparams)
        check(lib().cgx_problem_create(C.byref(self._h), model.kind, self.B, self.T, self.nx, model.nu, dptr(params),
                                       params.shape[-1], int(per), dptr(x0)))
        check(lib().cgx_problem_set_terminal_params(self._h, dptr(tparams), tparams.shape[-1]))
        lb, ub = getattr(model, "u_lb", None), getattr(model, "u_ub", None)
        if lb is not None and (np.isfinite(lb).any() or np.isfinite(ub).any()):
            self.set_u_bounds(lb, ub)

    def set_u_bounds(self, lb, ub):
        """set_u_lb / set_u_ub of the running model (action-base.hxx:117-140): [nu] shared or [B, nu]."""
        lb, ub = _f64(lb), _f64(ub)
        if lb.shape != ub.shape or lb.shape[-1] != self.nu_max or (lb.ndim == 2 and lb.shape[0] != self.B):
            raise CgxError(-1, "Invalid argument: lower bound has wrong dimension (it should be %d)" % self.nu_max)
        check(lib().cgx_problem_set_u_bounds(self._h, dptr(lb), dptr(ub), int(lb.ndim == 2)))

    def __del__(self):
        h = getattr(self, "_h", None)
        if h and lib is not None:  # module globals are already gone at interpreter shutdown
            lib().cgx_problem_destroy(h)
            self._h = None

    @property
    def x0(self):
        out = np.zeros((self.B, self.nx))
        check(lib().cgx_problem_get_x0(self._h, dptr(out)))
        return out

    @x0.setter
    def x0(self, v):
        v = _f64(v).reshape(self.B, self.nx)
        check(lib().cgx_problem_set_x0(self._h, dptr(v)))

    def _xs_us(self, xs, us):
        xs = _f64(xs).reshape(self.B, self.T + 1, self.nx) if np.size(xs) == self.B * (self.T + 1) * self.nx else None
        if xs is None:
            raise CgxError(-1, "Invalid argument: xs has wrong dimension (it should be %d)" % (self.T + 1))
        if np.size(us) != self.B * self.T * self.nu_max:
            raise CgxError(-1, "Invalid argument: us has wrong dimension (it should be %d)" % self.T)
        return xs, _f64(us).reshape(self.B, self.T, self.nu_max)

    def calc(self, xs, us):
        xs, us = self._xs_us(xs, us)
        costs = np.zeros(self.B)
        check(lib().cgx_problem_calc(self._h, dptr(xs), dptr(us), dptr(costs)))
        return costs

    def calcDiff(self, xs, us):
        xs, us = self._xs_us(xs, us)
        costs = np.zeros(self.B)
        check(lib().cgx_problem_calc_diff(self._h, dptr(xs), dptr(us), dptr(costs)))
        return costs

    def rollout(self, us):
        if np.size(us) != self.B * self.T * self.nu_max:
            raise CgxError(-1, "Invalid argument: us has wrong dimension (it should be %d)" % self.T)
        us = _f64(us).reshape(self.B, self.T, self.nu_max)
        xs = np.zeros((self.B, self.T + 1, self.nx))
        check(lib().cgx_problem_rollout(self._h, dptr(us), dptr(xs)))
        return xs

    def quasiStatic(self, xs):
        xs = _f64(xs).reshape(self.B, self.nx)
        us = np.zeros((self.B, self.nu_max))
        check(lib().cgx_problem_quasi_static(self._h, dptr(xs), dptr(us)))
        return us

    _DATA_SHAPES = {"cost": lambda n, m: (1,), "xnext": lambda n, m: (n,), "Fx": lambda n, m: (n, n),
                    "Fu": lambda n, m: (n, m), "Lx": lambda n, m: (n,), "Lu": lambda n, m: (m,),
                    "Lxx": lambda n, m: (n, n), "Lxu": lambda n, m: (n, m), "Luu": lambda n, m: (m, m)}

    def data(self, field, b, t):
        """Field of the ActionData of node t (t == T: terminalData) of problem b."""
        shp = (self.model.nr,) if field == "r" else self._DATA_SHAPES[field](self.ndx, self.nu_max)
        out = np.zeros(int(np.prod(shp)))
        check(lib().cgx_problem_get_data(self._h, field.encode(), b, t, dptr(out)))
        return out.reshape(shp, order="F")


class CallbackLogger(object):
    """CallbackLogger of the reference (bindings/python/crocoddyl/__init__.py:356-381): xs, us, fs, steps, iters, costs,
    u_regs, x_regs, stops, grads — for ONE problem of the batch (`problem`, default 0), filled from the device's
    per-iteration trace after the solve. steps / iters / costs / u_regs / x_regs / stops / grads hold one entry per
    iteration exactly like the reference's; xs and us are the solver's final iterate (what the reference's last call
    leaves there). The device loop does not return to the host between iterations, so `fs` holds the gaps of the LAST
    computeDirection only (one entry) instead of one entry per iteration. `records` keeps the raw trace rows."""

    def __init__(self, problem=0):
        self.problem = problem
        self.xs, self.us, self.fs = [], [], []
        self.steps, self.iters, self.costs, self.u_regs, self.x_regs, self.stops, self.grads = [], [], [], [], [], [], []
        self.records = []

    def __call__(self, rec):
        if rec["problem"] != self.problem:
            return
        self.records.append(dict(rec))
        self.steps.append(rec["steplength"])
        self.iters.append(rec["iter"])
        self.costs.append(rec["cost"])
        self.u_regs.append(rec["ureg"])
        self.x_regs.append(rec["xreg"])
        self.stops.append(rec["stop"])
        self.grads.append(-rec["d1"])  # -expectedImprovement()[1]

    def finalize(self, solver):
        b = self.problem
        self.xs = list(solver.get("xs")[b])
        self.us = list(solver.get("us")[b])
        self.fs = [list(solver.get("fs")[b])]


class CallbackVerbose(object):
    """Prints the table of src/core/utils/callbacks.cpp:17-67 for one problem of the batch."""

    def __init__(self, problem=0):
        self.problem = problem

    def __call__(self, rec):
        if rec["problem"] != self.problem:
            return
        if rec["iter"] % 10 == 0:
            print("iter \t cost \t      stop \t    grad \t  xreg \t      ureg \t step \t feas")
        print("%4d  %.5e  %.5e  %.5e  %.5e  %.5e  %.4f  %1d" % (rec["iter"], rec["cost"], rec["stop"], -rec["d1"],
                                                                rec["xreg"], rec["ureg"], rec["steplength"],
                                                                rec["is_feasible"]))


class SolverDDP(object):
    """SolverDDP(problem) (include/crocoddyl/core/solvers/ddp.hpp:48-302) over a batch of problems."""
    _kind = SOLVER_DDP
    _NT_PLUS = ("xs", "xs_try", "fs", "Vx", "Vxx")

    def __init__(self, problem):
        self.problem = problem
        self._h = C.c_void_p()
        check(lib().cgx_solver_create(C.byref(self._h), problem._h, self._kind))
        self._callbacks = []
        self._cb_keep = None

    def __del__(self):
        h = getattr(self, "_h", None)
        if h and lib is not None:  # module globals are already gone at interpreter shutdown
            lib().cgx_solver_destroy(h)
            self._h = None

    # ---- solve -------------------------------------------------------------------------------
    def _traj(self, xs, us):
        p = self.problem
        if xs is not None and np.size(xs) == 0:
            xs = None
        if us is not None and np.size(us) == 0:
            us = None
        if xs is not None:
            if np.size(xs) != p.B * (p.T + 1) * p.nx:
                raise CgxError(-1, "Warm start state has wrong dimension, got %d expecting %d" % (len(xs), p.T + 1))
            xs = _f64(xs).reshape(p.B, p.T + 1, p.nx)
        if us is not None:
            if np.size(us) != p.B * p.T * p.nu_max:
                raise CgxError(-1, "Warm start control has wrong dimension, got %d expecting %d" % (len(us), p.T))
            us = _f64(us).reshape(p.B, p.T, p.nu_max)
        return xs, us

    def solve(self, init_xs=None, init_us=None, maxiter=100, isFeasible=False, regInit=1e-9):
        """Returns the per-problem bool array the reference's solve() returns (a plain bool for B = 1)."""
        xs, us = self._traj(init_xs, init_us)
        conv = np.zeros(self.problem.B, dtype=np.uint8)
        reg = float("nan") if regInit is None else float(regInit)
        if self._callbacks:
            self._install_callbacks(maxiter)
        check(lib().cgx_solver_solve(self._h, dptr(xs), dptr(us), int(maxiter), int(bool(isFeasible)), reg,
                                     conv.ctypes.data_as(_lib._up)))
        conv = conv.astype(bool)
        self._finalize_callbacks()
        return bool(conv[0]) if self.problem.B == 1 else conv

    def solveFetch(self, init_xs, init_us, maxiter=100, isFeasible=False, regInit=1e-9):
        """solve() + the xs / us getters as one call (cgx_solver_solve_fetch): returns (converged, xs, us). On the
        thread-per-problem backend the host<->device copies of one slice of the batch overlap the solve of the others."""
        xs, us = self._traj(init_xs, init_us)
        p = self.problem
        conv = np.zeros(p.B, dtype=np.uint8)
        xs_out, us_out = np.empty_like(xs), np.empty_like(us)
        reg = float("nan") if regInit is None else float(regInit)
        if self._callbacks:
            self._install_callbacks(maxiter)
        check(lib().cgx_solver_solve_fetch(self._h, dptr(xs), dptr(us), int(maxiter), int(bool(isFeasible)), reg,
                                           conv.ctypes.data_as(_lib._up), dptr(xs_out), dptr(us_out)))
        self._finalize_callbacks()
        return conv.astype(bool), xs_out, us_out

    def _finalize_callbacks(self):
        for cb_ in self._callbacks:
            if hasattr(cb_, "finalize"):
                cb_.finalize(self)

    def setCandidate(self, xs_warm=None, us_warm=None, isFeasible=False):
        xs, us = self._traj(xs_warm, us_warm)
        check(lib().cgx_solver_set_candidate(self._h, dptr(xs), dptr(us), int(bool(isFeasible))))

    def computeDirection(self, recalc=True):
        st = np.zeros(self.problem.B, dtype=np.int32)
        check(lib().cgx_solver_compute_direction(self._h, int(bool(recalc)), st.ctypes.data_as(_lib._ip)))
        if st.any():
            raise CgxError(-1, "backward_error")

    def tryStep(self, stepLength=1.0):
        dV = np.zeros(self.problem.B)
        st = np.zeros(self.problem.B, dtype=np.int32)
        check(lib().cgx_solver_try_step(self._h, float(stepLength), dptr(dV), st.ctypes.data_as(_lib._ip)))
        if st.any():
            raise CgxError(-1, "forward_error")
        return dV

    def expectedImprovement(self):
        d = np.zeros((self.problem.B, 2))
        check(lib().cgx_solver_expected_improvement(self._h, dptr(d)))
        return d

    def stoppingCriteria(self):
        s = np.zeros(self.problem.B)
        check(lib().cgx_solver_stopping_criteria(self._h, dptr(s)))
        return s

    # ---- callbacks ---------------------------------------------------------------------------
    def setCallbacks(self, callbacks):
        self._callbacks = list(callbacks)

    def getCallbacks(self):
        return self._callbacks

    def _install_callbacks(self, maxiter):
        check(lib().cgx_solver_enable_trace(self._h, 1, max(int(maxiter), 1)))

        def tramp(_user, rec):
            r = rec.contents
            d = {k: getattr(r, k) for k, _ in _lib.IterRecord._fields_}
            for cb in self._callbacks:
                cb(d)

        self._cb_keep = _lib.CALLBACK(tramp)
        check(lib().cgx_solver_set_callback(self._h, self._cb_keep, None))

    def enableTrace(self, max_records=100):
        check(lib().cgx_solver_enable_trace(self._h, 1, int(max_records)))

    def trace(self, b=0):
        n = check(lib().cgx_solver_get_trace(self._h, b, None, 0))
        out = np.zeros((max(n, 1), _lib.NTRACE))
        check(lib().cgx_solver_get_trace(self._h, b, dptr(out), n))
        return out[:n]

    def setStorage(self, full):
        check(lib().cgx_solver_set_storage(self._h, int(bool(full))))

    # ---- accessors ---------------------------------------------------------------------------
    def _shape(self, name):
        n, m = self.problem.ndx, self.problem.nu_max
        return {"xs": (n,), "us": (m,), "xs_try": (n,), "us_try": (m,), "fs": (n,), "K": (m, n), "k": (m,),
                "Vx": (n,), "Vxx": (n, n), "Qxx": (n, n), "Qxu": (n, m), "Quu": (m, m), "Qx": (n,), "Qu": (m,),
                "Quu_inv": (m, m)}[name]

    def get(self, name):
        """[B, nt, *shape] array of a per-node field (matrices converted from column-major)."""
        shp = self._shape(name)
        p = self.problem
        nt = p.T + 1 if name in self._NT_PLUS else p.T
        out = np.zeros((p.B, nt, int(np.prod(shp))))
        check(lib().cgx_solver_get_all(self._h, name.encode(), dptr(out)))
        if len(shp) == 2:
            return np.transpose(out.reshape(p.B, nt, shp[1], shp[0]), (0, 1, 3, 2))
        return out

    def get_node(self, name, b, t):
        shp = self._shape(name)
        out = np.zeros(int(np.prod(shp)))
        check(lib().cgx_solver_get(self._h, name.encode(), b, t, dptr(out)))
        return out.reshape(shp, order="F")

    def scalars(self):
        out = np.zeros((self.problem.B, _lib.NSCALARS))
        check(lib().cgx_solver_get_scalars(self._h, dptr(out)))
        return {k: out[:, i].copy() for i, k in enumerate(_lib.SCALARS)}

    def timing(self):
        ms = np.zeros(3)
        n = C.c_int(0)
        check(lib().cgx_solver_last_timing(self._h, dptr(ms), C.byref(n)))
        return {"call_ms": ms[0], "kernel_ms": ms[1], "launches": n.value}

    def _param(self, name):
        v = C.c_double(0)
        check(lib().cgx_solver_get_param(self._h, name.encode(), C.byref(v)))
        return v.value

    def _set_param(self, name, v):
        check(lib().cgx_solver_set_param(self._h, name.encode(), float(v)))

    @property
    def alphas(self):
        a = np.zeros(16)
        n = C.c_int(0)
        check(lib().cgx_solver_get_alphas(self._h, dptr(a), C.byref(n)))
        return a[:n.value].copy()

    @alphas.setter
    def alphas(self, v):
        v = _f64(v)
        check(lib().cgx_solver_set_alphas(self._h, dptr(v), v.size))


def _field_prop(name):
    return property(lambda self: self.get(name))


def _scalar_prop(name):
    return property(lambda self: self.scalars()[name])


def _param_prop(cname):
    return property(lambda self: self._param(cname), lambda self, v: self._set_param(cname, v))


for _n in ("K", "k", "Vx", "Vxx", "Qxx", "Qxu", "Quu", "Qx", "Qu", "fs"):
    setattr(SolverDDP, _n, _field_prop(_n))
SolverDDP.xs = property(lambda self: self.get("xs"),
                        lambda self, v: check(lib().cgx_solver_set_xs(self._h, dptr(self._traj(v, None)[0]))))
SolverDDP.us = property(lambda self: self.get("us"),
                        lambda self, v: check(lib().cgx_solver_set_us(self._h, dptr(self._traj(None, v)[1]))))
for _py, _c in (("cost", "cost"), ("stop", "stop"), ("stepLength", "steplength"), ("dV", "dV"), ("dVexp", "dVexp"),
                ("iter", "iter"), ("isFeasible", "is_feasible")):
    setattr(SolverDDP, _py, _scalar_prop(_c))
SolverDDP.x_reg = property(lambda self: self.scalars()["xreg"], lambda self, v: self._set_param("xreg", v))
SolverDDP.u_reg = property(lambda self: self.scalars()["ureg"], lambda self, v: self._set_param("ureg", v))
SolverDDP.d = property(lambda self: np.stack([self.scalars()["d0"], self.scalars()["d1"]], axis=1))
for _py, _c in (("th_acceptStep", "th_acceptstep"), ("th_stop", "th_stop"), ("th_grad", "th_grad"),
                ("th_stepDec", "th_stepdec"), ("th_stepInc", "th_stepinc"), ("regFactor", "regfactor"),
                ("regMin", "regmin"), ("regMax", "regmax")):
    setattr(SolverDDP, _py, _param_prop(_c))


class SolverFDDP(SolverDDP):
    """SolverFDDP(problem) (include/crocoddyl/core/solvers/fddp.hpp:50-100)."""
    _kind = SOLVER_FDDP
    th_acceptNegStep = _param_prop("th_acceptnegstep")


class SolverBoxDDP(SolverDDP):
    """SolverBoxDDP(problem) (include/crocoddyl/core/solvers/box-ddp.hpp, src/core/solvers/box-ddp.cpp): control-limited
    DDP — BoxQP in computeGains, clamp in forwardPass; reads u_lb / u_ub of the running model."""
    _kind = 2
    Quu_inv = _field_prop("Quu_inv")


class SolverBoxFDDP(SolverFDDP):
    """SolverBoxFDDP(problem) (box-fddp.hpp, src/core/solvers/box-fddp.cpp)."""
    _kind = 3
    Quu_inv = _field_prop("Quu_inv")


def device_count():
    return lib().cgx_device_count()


def set_device(i):
    check(lib().cgx_set_device(int(i)))


def probe_fp64_tflops():
    v = C.c_double(0)
    check(lib().cgx_probe_fp64_tflops(C.byref(v)))
    return v.value
