"""Fixed-base revolute-chain arm models for the batched DDP/FDDP path.

Mirrors, for this path only, the objects the reference's arm benchmark builds (benchmark/factory/arm.hpp:60-90):
IntegratedActionModelEuler(DifferentialActionModelFreeFwdDynamics(StateMultibody, ActuationModelFull,
CostModelSum{gripperPose: CostModelFramePlacement, xReg: CostModelState, uReg: CostModelControl}), dt) for the
running nodes and the SAME differential model with dt = 0 for the terminal node. The robot is a parameter table
instead of a URDF (no Pinocchio / example-robot-data on this path): per joint a unit axis, the placement in the
parent joint frame, and the body inertia (mass, COM lever, rotational inertia about the COM).

Options of the same differential model: `armature` (free-fwddyn.hxx:66-74,108-114, as
examples/arm_manipulation.py:53 sets it), `xWeights` = ActivationModelWeightedQuad on the state cost
(weighted-quadratic.hpp), `goal` = "placement" | "translation" | "rotation" for CostModelFramePlacement /
CostModelFrameTranslation / CostModelFrameRotation on the gripper frame.

Parameter block (doubles) = 36 + 28 nv, see include/crocoddyl_b200.h "ARM":
  [0] dt [1..3] gravity [4] w_gripperPose [5] w_xReg [6] w_uReg [7] frame parent joint
  [8..19] frame placement R(9 row-major) p(3)   [20..31] Mref R(9) p(3)   [32] goal kind (0/1/2)   [33..35] reserved
  per joint: axis(3) placement R(9) p(3) mass com(3) Ixx Ixy Ixz Iyy Iyz Izz armature w_q w_v
"""
import numpy as np

from .base import ActionModelAbstract

MODEL_ARM = 2


ARM_HEADER, ARM_PER_JOINT = 36, 28
GOAL_KINDS = {"placement": 0, "translation": 1, "rotation": 2}


class RevoluteChain(object):
    """Kinematic/inertial description of a fixed-base serial chain of 1-DoF revolute joints."""

    def __init__(self, axes, placements_R, placements_p, masses, coms, inertias, gravity=(0., 0., -9.81)):
        self.axes = np.asarray(axes, dtype=np.float64)
        self.nv = self.axes.shape[0]
        self.R = np.asarray(placements_R, dtype=np.float64).reshape(self.nv, 3, 3)
        self.p = np.asarray(placements_p, dtype=np.float64).reshape(self.nv, 3)
        self.masses = np.asarray(masses, dtype=np.float64).reshape(self.nv)
        self.coms = np.asarray(coms, dtype=np.float64).reshape(self.nv, 3)
        self.inertias = np.asarray(inertias, dtype=np.float64).reshape(self.nv, 3, 3)
        self.gravity = np.asarray(gravity, dtype=np.float64)
        nrm = np.linalg.norm(self.axes, axis=1)
        if np.any(np.abs(nrm - 1.) > 1e-12):
            raise ValueError("joint axes must be unit vectors")


def synthetic_arm(nv=7):
    """The synthetic 7-DoF arm of SURVEY 8d / BASELINE configs C4, C5: axes alternate Z,Y,Z,Y,...; every joint sits
    (0,0,0.3) above its parent joint; link mass 2 kg, COM (0,0,0.15), inertia diag(0.02,0.02,0.005) about the COM."""
    axes = [(0., 0., 1.) if i % 2 == 0 else (0., 1., 0.) for i in range(nv)]
    R = [np.eye(3)] * nv
    p = [(0., 0., 0.3)] * nv
    return RevoluteChain(axes, R, p, [2.0] * nv, [(0., 0., 0.15)] * nv, [np.diag([0.02, 0.02, 0.005])] * nv)


class ArmActionModel(ActionModelAbstract):
    """IntegratedActionModelEuler(DifferentialActionModelFreeFwdDynamics(...), dt) (euler.hxx, free-fwddyn.hxx) with
    the cost sum of benchmark/factory/arm.hpp:62-74. `terminal_model` is the same differential model with dt = 0."""
    kind = MODEL_ARM

    def __init__(self, chain, dt=1e-3, frame_parent=None, frame_R=None, frame_p=(0., 0., 0.1), Mref_R=None,
                 Mref_p=(0., 0., 0.4), w_gripperPose=1., w_xReg=1e-4, w_uReg=1e-4, armature=None, xWeights=None,
                 goal="placement", _terminal_of=None):
        self.chain = chain
        self.nv = chain.nv
        self._init_abstract(2 * chain.nv, chain.nv, 6 + 2 * chain.nv + chain.nv, nq=chain.nv)
        self.dt = float(dt)
        self.frame_parent = chain.nv - 1 if frame_parent is None else int(frame_parent)
        self.frame_R = np.eye(3) if frame_R is None else np.asarray(frame_R, dtype=np.float64)
        self.frame_p = np.asarray(frame_p, dtype=np.float64)
        self.Mref_R = np.eye(3) if Mref_R is None else np.asarray(Mref_R, dtype=np.float64)
        self.Mref_p = np.asarray(Mref_p, dtype=np.float64)
        self.w_gripperPose, self.w_xReg, self.w_uReg = float(w_gripperPose), float(w_xReg), float(w_uReg)
        self.armature = np.zeros(chain.nv) if armature is None else np.asarray(armature, dtype=np.float64).reshape(chain.nv)
        self.xWeights = (np.ones(2 * chain.nv) if xWeights is None
                         else np.asarray(xWeights, dtype=np.float64).reshape(2 * chain.nv))
        if goal not in GOAL_KINDS:
            raise ValueError("goal must be one of %s" % sorted(GOAL_KINDS))
        self.goal = goal
        if _terminal_of is None:
            self.terminal_of = ArmActionModel(chain, 0., self.frame_parent, self.frame_R, self.frame_p, self.Mref_R,
                                              self.Mref_p, w_gripperPose, w_xReg, w_uReg, self.armature, self.xWeights,
                                              goal, _terminal_of=self)
        else:
            self.terminal_of = None

    @property
    def terminal_model(self):
        return self.terminal_of

    def params(self):
        c = self.chain
        out = [self.dt] + list(c.gravity) + [self.w_gripperPose, self.w_xReg, self.w_uReg, float(self.frame_parent)]
        out += list(self.frame_R.ravel()) + list(self.frame_p) + list(self.Mref_R.ravel()) + list(self.Mref_p)
        out += [float(GOAL_KINDS[self.goal]), 0., 0., 0.]
        for i in range(c.nv):
            I = c.inertias[i]
            out += list(c.axes[i]) + list(c.R[i].ravel()) + list(c.p[i]) + [c.masses[i]] + list(c.coms[i])
            out += [I[0, 0], I[0, 1], I[0, 2], I[1, 1], I[1, 2], I[2, 2]]
            out += [self.armature[i], self.xWeights[i], self.xWeights[c.nv + i]]
        p = np.asarray(out, dtype=np.float64)
        assert p.size == ARM_HEADER + ARM_PER_JOINT * c.nv
        return p
