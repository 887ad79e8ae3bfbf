"""Node plugin interface of the reference for the batched device path: StateVector, ActionDataAbstract,
ActionModelAbstract (include/crocoddyl/core/action-base.hpp:34-142, core/states/euclidean.hxx:28-61)."""
import numpy as np

from ._lib import CgxError


def _f64(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if shape is not None:
        a = a.reshape(shape)
    return a


class StateVector(object):
    """StateVector(nx) (include/crocoddyl/core/states/euclidean.hxx:28-61): the state of every model on this path
    (StateMultibody of a fixed-base revolute chain is the same vector space, multibody/states/multibody.hxx:69-70,89-90)."""

    def __init__(self, nx, nq=None):
        self.nx = self.ndx = int(nx)
        self.nq = self.nx // 2 if nq is None else int(nq)
        self.nv = self.nx - self.nq

    def zero(self):
        return np.zeros(self.nx)

    def diff(self, x0, x1):
        return np.asarray(x1, dtype=np.float64) - np.asarray(x0, dtype=np.float64)

    def integrate(self, x, dx):
        return np.asarray(x, dtype=np.float64) + np.asarray(dx, dtype=np.float64)


class ActionDataAbstract(object):
    """ActionDataAbstract (include/crocoddyl/core/action-base.hpp:101-142): cost, xnext, r, Fx, Fu, Lx, Lu, Lxx, Lxu, Luu.
    Batched models carry a leading batch dimension on every field. `r` (the cost residual) is filled where the
    reference's model writes it: ActionModelUnicycle::calc (unicycle.hxx:39-40). ActionModelLQR has nr = 0, and the
    reference's cost sum does not write the action data's r for the integrated arm model, so it stays zero there."""

    def __init__(self, model):
        n, m = model.ndx, model.nu
        self.cost = 0.
        self.xnext = np.zeros(model.nx)
        self.r = np.zeros(model.nr)
        self.Fx, self.Fu = np.zeros((n, n)), np.zeros((n, m))
        self.Lx, self.Lu = np.zeros(n), np.zeros(m)
        self.Lxx, self.Lxu, self.Luu = np.zeros((n, n)), np.zeros((n, m)), np.zeros((m, m))


class ActionModelAbstract(object):
    """The node plugin interface of the reference (include/crocoddyl/core/action-base.hpp:34-98, action-base.hxx) for the
    model families the device path implements: calc(data, x, u), calcDiff(data, x, u), createData(), quasiStatic(data, x),
    nu / nr / state, u_lb / u_ub / has_control_limits. calc / calcDiff evaluate ONE node on the GPU (a one-node
    ShootingProblem through the C ABI); x may be [nx] or a batch [B, nx]. u = None is the x-only overload (unone = 0),
    i.e. how the reference evaluates a terminal node (action-base.hxx:27-37)."""
    kind = None

    def _init_abstract(self, nx, nu, nr, nq=None):
        self.nx = self.ndx = int(nx)
        self.nu = int(nu)
        self.nr = int(nr)
        self.state = StateVector(nx, nq)
        self.unone = np.zeros(self.nu)
        self._u_lb = np.full(self.nu, -np.inf)
        self._u_ub = np.full(self.nu, np.inf)
        self._node = None

    # control limits (action-base.hxx:107-144)
    @property
    def u_lb(self):
        return self._u_lb

    @u_lb.setter
    def u_lb(self, v):
        v = _f64(v)
        if v.shape[-1] != self.nu:
            raise CgxError(-1, "Invalid argument: lower bound has wrong dimension (it should be %d)" % self.nu)
        self._u_lb = v

    @property
    def u_ub(self):
        return self._u_ub

    @u_ub.setter
    def u_ub(self, v):
        v = _f64(v)
        if v.shape[-1] != self.nu:
            raise CgxError(-1, "Invalid argument: upper bound has wrong dimension (it should be %d)" % self.nu)
        self._u_ub = v

    @property
    def has_control_limits(self):
        return bool(np.isfinite(self._u_lb).any() and np.isfinite(self._u_ub).any())

    def createData(self):
        return ActionDataAbstract(self)

    def checkData(self, data):
        return isinstance(data, ActionDataAbstract)

    def _node_problem(self, x):
        """A one-node ShootingProblem (running = terminal = this model) on the device for the batch size of x."""
        x = _f64(x)
        xb = x[None, :] if x.ndim == 1 else x
        if xb.shape[1] != self.nx:
            raise CgxError(-1, "Invalid argument: x has wrong dimension (it should be %d)" % self.nx)
        key = (xb.shape[0], self.params().tobytes())
        if self._node is None or self._node[0] != key:
            from . import ShootingProblem  # the package imports this module first
            self._node = (key, ShootingProblem(xb, [self], self))
        return self._node[1], xb, x.ndim == 1

    def _fill(self, data, prob, t, one, fields):
        for f in fields:
            v = np.stack([prob.data(f, b, t) for b in range(prob.B)])
            if f == "cost":
                v = v[:, 0]
            setattr(data, f, v[0] if one else v)

    def _eval(self, data, x, u, diff):
        prob, xb, one = self._node_problem(x)
        B = xb.shape[0]
        t = 0 if u is not None else 1  # x-only overload = the terminal node of the one-node problem
        if u is not None:
            ub = _f64(u).reshape(B, -1)
            if ub.shape[1] != self.nu:
                raise CgxError(-1, "Invalid argument: u has wrong dimension (it should be %d)" % self.nu)
        else:
            ub = np.zeros((B, self.nu))
        xs = np.repeat(xb[:, None, :], 2, 1)
        prob.calc(xs, ub[:, None, :])
        self._fill(data, prob, t, one, ("cost", "xnext") + (("r",) if getattr(self, "writes_residual", False) else ()))
        if diff:
            prob.calcDiff(xs, ub[:, None, :])
            self._fill(data, prob, t, one, ("Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu"))

    def calc(self, data, x, u=None):
        self._eval(data, x, u, False)

    def calcDiff(self, data, x, u=None):
        """calcDiff (the reference expects calc to have been called; here both are evaluated so that `data` is complete)."""
        self._eval(data, x, u, True)

    def quasiStatic(self, data, x, maxiter=100, tol=1e-9):
        """ActionModelAbstract::quasiStatic (action-base.hxx:39-83); closed form for the arm (free-fwddyn.hxx:146-169)."""
        prob, xb, one = self._node_problem(x)
        u = prob.quasiStatic(xb)
        return u[0] if one else u


