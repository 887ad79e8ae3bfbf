"""IntegratedActionModelEuler over a differential action model (include/crocoddyl/core/integrator/euler.hxx:14-131) for
the two differential models of this path, each mapped onto the node model whose kernels already exist:

  IntegratedActionModelEuler(DifferentialActionModelLQR(nq, nu), dt)   -> an ActionModelLQR with the discrete matrices
      of the semi-implicit Euler step (runs on the LQR kernels; SURVEY 8 f3 "DifferentialActionModelLQR on the same
      kernels"): v' = v + dt a, q' = q + dt v' with a = Fq q + Fv v + Fu u (+ f0) is linear in (x, u);
  IntegratedActionModelEuler(DifferentialActionModelFreeFwdDynamics(chain, ...), dt) -> the ArmActionModel.
"""
import numpy as np

from .arm import ArmActionModel

MODEL_LQR = 0


class DifferentialActionModelLQR(object):
    """DifferentialActionModelLQR(nq, nu, driftFree=True) (include/crocoddyl/core/actions/diff-lqr.hxx:14-88):
    a = Fq q + Fv v + Fu u (+ f0), cost = 1/2 [x,u]^T [Lxx Lxu; Lxu^T Luu] [x,u] + lx^T x + lu^T u. Defaults as the
    reference's constructor (identities and ones). Matrices may carry a leading batch dimension."""

    def __init__(self, nq, nu, driftFree=True):
        self.nq, self.nv, self.nu = int(nq), int(nq), int(nu)
        self.nx = 2 * self.nq
        self.driftFree = bool(driftFree)
        self.Fq = np.eye(nq)
        self.Fv = np.eye(nq)
        self.Fu = np.eye(nq, nu)
        self.f0 = np.ones(nq)
        self.Lxx = np.eye(self.nx)
        self.Lxu = np.eye(self.nx, nu)
        self.Luu = np.eye(nu)
        self.lx = np.ones(self.nx)
        self.lu = np.ones(nu)

    def calc(self, x, u):
        """(xout, cost) of diff-lqr.hxx:48-57 for one state/control (host arithmetic on the model's own matrices: this
        is the differential model's definition, the solve path never calls it)."""
        x, u = np.asarray(x, dtype=np.float64), np.asarray(u, dtype=np.float64)
        q, v = x[:self.nq], x[self.nq:]
        xout = self.Fq @ q + self.Fv @ v + self.Fu @ u
        if not self.driftFree:
            xout = xout + self.f0
        cost = 0.5 * x @ (self.Lxx @ x) + 0.5 * u @ (self.Luu @ u) + x @ (self.Lxu @ u) + self.lx @ x + self.lu @ u
        return xout, cost


class DifferentialActionModelFreeFwdDynamics(object):
    """DifferentialActionModelFreeFwdDynamics(state, actuation = full, costs = {gripperPose, xReg, uReg})
    (include/crocoddyl/multibody/actions/free-fwddyn.hxx:14-169) for a fixed-base revolute chain; the arguments are
    the ones ArmActionModel takes (the chain and the cost sum of benchmark/factory/arm.hpp:62-74)."""

    def __init__(self, chain, **arm_kwargs):
        self.chain = chain
        self.arm_kwargs = dict(arm_kwargs)
        self.nq = self.nv = self.nu = chain.nv
        self.nx = 2 * chain.nv


def _euler_lqr_matrices(d, dt):
    """Discrete (Fx, Fu, f0, Lxx, Lxu, Luu, lx, lu) of euler.hxx:65-74,100-130 over diff-lqr.hxx:76-84, in the
    reference's operation order (da * dt^2, da * dt, + dt on the top-right diagonal, + identity)."""
    nq, nu, nx = d.nq, d.nu, d.nx

    def lead(a, nd):
        a = np.asarray(a, dtype=np.float64)
        return a.shape[:-nd] if a.ndim > nd else ()

    batch = ()
    for name, nd in (("Fq", 2), ("Fv", 2), ("Fu", 2), ("f0", 1), ("Lxx", 2), ("Lxu", 2), ("Luu", 2), ("lx", 1), ("lu", 1)):
        b = lead(getattr(d, name), nd)
        if b:
            if batch and b != batch:
                raise ValueError("inconsistent batch dimension in DifferentialActionModelLQR." + name)
            batch = b
    g = lambda name: np.asarray(getattr(d, name), dtype=np.float64)  # noqa: E731
    Fq, Fv, Fu, f0 = g("Fq"), g("Fv"), g("Fu"), g("f0")
    Fx = np.zeros(batch + (nx, nx))
    Fud = np.zeros(batch + (nx, nu))
    f0d = np.zeros(batch + (nx,))
    if dt != 0.:
        dt2 = dt * dt
        Fx[..., :nq, :nq] = Fq * dt2
        Fx[..., :nq, nq:] = Fv * dt2
        Fx[..., nq:, :nq] = Fq * dt
        Fx[..., nq:, nq:] = Fv * dt
        idx = np.arange(nq)
        Fx[..., idx, nq + idx] += dt
        Fx[..., np.arange(nx), np.arange(nx)] += 1.
        Fud[..., :nq, :] = Fu * dt2
        Fud[..., nq:, :] = Fu * dt
        f0d[..., :nq] = f0 * dt2
        f0d[..., nq:] = f0 * dt
        sc = dt
    else:
        Fx[..., np.arange(nx), np.arange(nx)] = 1.
        sc = 1.
    return dict(Fx=Fx, Fu=Fud, f0=f0d, Lxx=sc * g("Lxx"), Lxu=sc * g("Lxu"), Luu=sc * g("Luu"), lx=sc * g("lx"),
                lu=sc * g("lu"))


def make_integrated_euler(ActionModelLQR):
    """Builds the IntegratedActionModelEuler factory once ActionModelLQR exists (package __init__)."""

    class _EulerDiffLQR(ActionModelLQR):
        """IntegratedActionModelEuler(DifferentialActionModelLQR, dt) as an ActionModelLQR whose matrices are derived
        from the differential model every time the parameter block is packed (so later edits of the differential
        model's matrices count, like the reference's shared_ptr aliasing)."""

        def __init__(self, differential, timeStep):
            ActionModelLQR.__init__(self, differential.nx, differential.nu, True)
            self.differential = differential
            self.timeStep = float(timeStep)
            self._sync()

        def _sync(self):
            # with dt = 0 the step is the identity: no drift term whatever the differential model says
            self.driftFree = self.differential.driftFree or self.timeStep == 0.
            for k, v in _euler_lqr_matrices(self.differential, self.timeStep).items():
                setattr(self, k, v)

        def params(self):
            self._sync()
            return ActionModelLQR.params(self)

    def IntegratedActionModelEuler(differential, timeStep=1e-3, withCostResidual=True):
        """IntegratedActionModelEuler(model, timeStep = 1e-3, withCostResidual = true) (euler.hxx:14-34; a negative
        time step falls back to 1e-3 with a warning there, here too)."""
        dt = float(timeStep)
        if dt < 0.:
            dt = 1e-3
        if isinstance(differential, DifferentialActionModelLQR):
            return _EulerDiffLQR(differential, dt)
        if isinstance(differential, DifferentialActionModelFreeFwdDynamics):
            model = ArmActionModel(differential.chain, dt, **differential.arm_kwargs)
            model.differential = differential
            return model
        raise TypeError("IntegratedActionModelEuler: unsupported differential model %r (DifferentialActionModelLQR and "
                        "DifferentialActionModelFreeFwdDynamics are built)" % type(differential).__name__)

    return IntegratedActionModelEuler
