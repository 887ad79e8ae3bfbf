"""Third-party pins for the arm rows (the reference holds no vectors for Pinocchio's algorithms, and
Pinocchio is absent from /root/reference). Two anchors that owe nothing to this repository's own rigid-body code:

  * scipy.linalg.logm of the 4x4 homogeneous matrix for log6 (pinocchio/spatial/log.hxx is the published closed form of
    that matrix logarithm), including theta -> 0 and theta -> pi;
  * the textbook two-link planar arm (Spong, Hutchinson, Vidyasagar, "Robot Modeling and Control", eq. for the planar
    elbow manipulator): M(q), C(q, v), g(q) in closed form for CRBA / RNEA / ABA, and complex-step differentiation of
    that closed form (exact to rounding) for the ABA derivatives.

Both are applied to the oracle (oracle/rbd.hpp through orc_arm_op) AND to the product's own arithmetic
(crocoddyl_b200/csrc/rigid_body.cuh instantiated on the host, tests/host_rb): node cost through log6 of the frame
placement, qdd, and the Fx / Fu blocks of the semi-implicit Euler step."""
import numpy as np
import pytest
from scipy.linalg import expm, logm

import crocoddyl_b200 as cb
import oracle_lib as ol
from test_rigid_body_host_cpu import _eval, host_rb  # noqa: F401  (fixture)

G = 9.81


# ---------------------------------------------------------------- log6 against scipy.linalg.logm
def _hat(w):
    return np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]])


def _twist(xi):
    T = np.zeros((4, 4))
    T[:3, :3] = _hat(xi[3:])
    T[:3, 3] = xi[:3]
    return T


def _logm6(M):
    L = np.real(logm(M))
    return np.array([L[0, 3], L[1, 3], L[2, 3], L[2, 1], L[0, 2], L[1, 0]])


def _se3_of(theta, axis, p):
    axis = np.asarray(axis, dtype=float) / np.linalg.norm(axis)
    M = np.eye(4)
    M[:3, :3] = expm(_hat(theta * axis))
    M[:3, 3] = p
    return M


_ANGLES = [0.7, 2.9, np.pi - 1e-3, np.pi - 1e-5, 1e-2, 1e-5]


@pytest.mark.parametrize("theta", _ANGLES)
def test_oracle_log6_equals_scipy_logm(theta):
    P = cb.ArmActionModel(cb.synthetic_arm(7)).params()
    rng = np.random.default_rng(int(theta * 1e6) % 1000)
    for _ in range(4):
        M = _se3_of(theta, rng.standard_normal(3), rng.standard_normal(3))
        got = ol.arm_op(P, "log6", np.concatenate([M[:3, :3].ravel(), M[:3, 3]]), 6)
        want = _logm6(M)
        # logm (Schur + inverse scaling and squaring) itself is good to ~1e-10 near pi; acos near 0 costs eps/theta^2
        tol = 1e-9 if theta > 1e-3 else 1e-6
        np.testing.assert_allclose(got, want, rtol=tol, atol=tol * 1e-2)


def test_oracle_log6_of_pure_translation_and_identity():
    P = cb.ArmActionModel(cb.synthetic_arm(7)).params()
    for p in ([0., 0., 0.], [0.3, -1.2, 2.0]):
        M = _se3_of(0., [1, 0, 0], p)
        got = ol.arm_op(P, "log6", np.concatenate([M[:3, :3].ravel(), M[:3, 3]]), 6)
        np.testing.assert_allclose(got, np.concatenate([p, np.zeros(3)]), atol=1e-15)


def _fk(model, q):
    """Frame placement by plain 4x4 products: prod_i [R_i p_i] Rot(axis_i, q_i), then the frame offset."""
    c = model.chain
    M = np.eye(4)
    for i in range(model.frame_parent + 1):
        Pl = np.eye(4)
        Pl[:3, :3], Pl[:3, 3] = c.R[i], c.p[i]
        M = M @ Pl @ _se3_of(q[i], c.axes[i], np.zeros(3))
    F = np.eye(4)
    F[:3, :3], F[:3, 3] = model.frame_R, model.frame_p
    return M @ F


@pytest.mark.parametrize("nv,seed", [(7, 0), (7, 1), (4, 2), (2, 3)])
def test_product_goal_cost_equals_scipy_logm_of_the_frame_error(host_rb, nv, seed):  # noqa: F811
    """The terminal node with only the gripper-pose weight: cost = 0.5 |log6(Mref^-1 oMf(q))|^2
    (frame-placement.hxx:29-38, cost-base activation 0.5 r^T r), the logarithm taken by scipy on the 4x4 matrix and
    the forward kinematics by plain matrix products."""
    from test_oracle_arm_cpu import _random_chain
    rng = np.random.default_rng(seed)
    chain = _random_chain(nv, seed)
    Q, _r = np.linalg.qr(rng.standard_normal((3, 3)))
    if np.linalg.det(Q) < 0:
        Q[:, 0] *= -1
    model = cb.ArmActionModel(chain, 1e-3, frame_p=(0.05, -0.02, 0.1), Mref_R=Q, Mref_p=rng.standard_normal(3) * 0.3,
                              w_gripperPose=1., w_xReg=0., w_uReg=0.)
    Mref = np.eye(4)
    Mref[:3, :3], Mref[:3, 3] = model.Mref_R, model.Mref_p
    for _ in range(4):
        q = rng.uniform(-2., 2., nv)
        x = np.concatenate([q, rng.standard_normal(nv)])
        r = _logm6(np.linalg.inv(Mref) @ _fk(model, q))
        want = 0.5 * r @ r
        got = _eval(host_rb, nv, "c", model.terminal_model.params(), True, x, None)[0]
        np.testing.assert_allclose(got, want, rtol=1e-9)
        ob = ol.OracleBatch("arm", "fddp", 1, 2 * nv, nv, model.params(), x[None])
        ob.set_candidate(np.stack([x, x])[None], np.zeros((1, 1, nv)), False)
        ob.calc()
        np.testing.assert_allclose(ob.get("node_cost", 0, 1)[0], want, rtol=1e-9)


# ---------------------------------------------------------------- two-link planar arm, closed form
class Planar2R(object):
    """Both joints about +Y, link 1 of length l1 along +Z at q = 0, gravity -Z. Angles from the vertical."""

    def __init__(self, m1=1.3, m2=0.9, l1=0.45, lc1=0.2, lc2=0.25, I1=0.031, I2=0.017):
        self.m1, self.m2, self.l1, self.lc1, self.lc2, self.I1, self.I2 = m1, m2, l1, lc1, lc2, I1, I2

    def chain(self):
        I = lambda Iyy: np.diag([0.011, Iyy, 0.007])  # only Iyy enters planar motion
        return cb.RevoluteChain([(0., 1., 0.)] * 2, [np.eye(3)] * 2, [(0., 0., 0.), (0., 0., self.l1)],
                                [self.m1, self.m2], [(0., 0., self.lc1), (0., 0., self.lc2)], [I(self.I1), I(self.I2)])

    def M(self, q):
        c2 = np.cos(q[1])
        a = self.m2 * self.l1 * self.lc2
        m11 = self.m1 * self.lc1 ** 2 + self.m2 * (self.l1 ** 2 + self.lc2 ** 2) + 2 * a * c2 + self.I1 + self.I2
        m12 = self.m2 * self.lc2 ** 2 + a * c2 + self.I2
        m22 = self.m2 * self.lc2 ** 2 + self.I2
        return np.array([[m11, m12], [m12, m22]])

    def bias(self, q, v):
        h = -self.m2 * self.l1 * self.lc2 * np.sin(q[1])
        cor = np.array([h * v[1] * v[0] + h * (v[0] + v[1]) * v[1], -h * v[0] * v[0]])
        grav = np.array([-(self.m1 * self.lc1 + self.m2 * self.l1) * G * np.sin(q[0]) - self.m2 * self.lc2 * G * np.sin(q[0] + q[1]),
                         -self.m2 * self.lc2 * G * np.sin(q[0] + q[1])])
        return cor + grav

    def aba(self, q, v, tau):
        M = self.M(q)
        r = tau - self.bias(q, v)
        det = M[0, 0] * M[1, 1] - M[0, 1] * M[1, 0]
        return np.array([M[1, 1] * r[0] - M[0, 1] * r[1], -M[1, 0] * r[0] + M[0, 0] * r[1]]) / det

    def aba_derivatives(self, q, v, tau):
        """Complex-step differentiation of the closed form (no subtraction: exact to rounding)."""
        h = 1e-30
        z = np.concatenate([q, v, tau]).astype(complex)
        J = np.zeros((2, 6))
        for k in range(6):
            zz = z.copy()
            zz[k] += 1j * h
            J[:, k] = np.imag(self.aba(zz[:2], zz[2:4], zz[4:])) / h
        return J[:, :2], J[:, 2:4], J[:, 4:]


_STATES = [((0.3, -0.8), (0.0, 0.0), (0.0, 0.0)), ((-1.1, 2.0), (0.7, -1.9), (0.5, -0.2)),
           ((2.5, 0.0), (-1.0, 0.4), (-0.2, 1.4)), ((0.0, np.pi / 2), (3.0, 3.0), (0.0, 0.0))]


@pytest.mark.parametrize("state", _STATES)
def test_oracle_crba_rnea_aba_equal_planar_2r_closed_form(state):
    arm = Planar2R()
    P = cb.ArmActionModel(arm.chain(), 1e-3).params()
    q, v, tau = [np.array(s, dtype=float) for s in state]
    M = ol.arm_op(P, "crba", q, 4).reshape(2, 2, order="F")
    np.testing.assert_allclose(M, arm.M(q), rtol=1e-13, atol=1e-15)
    a = arm.aba(q, v, tau)
    np.testing.assert_allclose(ol.arm_op(P, "aba", np.concatenate([q, v, tau]), 2), a, rtol=1e-12, atol=1e-13)
    np.testing.assert_allclose(ol.arm_op(P, "rnea", np.concatenate([q, v, a]), 2), tau, rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(ol.arm_op(P, "rnea", np.concatenate([q, v, 0 * a]), 2), arm.bias(q, v), rtol=1e-12, atol=1e-13)


@pytest.mark.parametrize("state", _STATES)
def test_oracle_aba_derivatives_equal_complex_step_of_the_closed_form(state):
    arm = Planar2R()
    P = cb.ArmActionModel(arm.chain(), 1e-3).params()
    q, v, tau = [np.array(s, dtype=float) for s in state]
    out = ol.arm_op(P, "aba_derivatives", np.concatenate([q, v, tau]), 12)
    dq, dv, Minv = [out[i * 4:(i + 1) * 4].reshape(2, 2, order="F") for i in range(3)]
    wq, wv, wt = arm.aba_derivatives(q, v, tau)
    np.testing.assert_allclose(dq, wq, rtol=1e-11, atol=1e-11)
    np.testing.assert_allclose(dv, wv, rtol=1e-11, atol=1e-11)
    np.testing.assert_allclose(Minv, wt, rtol=1e-11, atol=1e-12)
    np.testing.assert_allclose(Minv, np.linalg.inv(arm.M(q)), rtol=1e-11, atol=1e-12)


@pytest.mark.parametrize("state", _STATES)
def test_product_node_equals_planar_2r_closed_form(host_rb, state):  # noqa: F811
    """The product's calc / calcDiff (rigid_body.cuh on the host): qdd, the semi-implicit Euler step
    (euler.hxx:44-63: v' = v + dt a, q' = q + dt v') and its Jacobians (euler.hxx:80-107) from the closed form."""
    arm = Planar2R()
    dt = 5e-3
    model = cb.ArmActionModel(arm.chain(), dt)
    q, v, tau = [np.array(s, dtype=float) for s in state]
    x = np.concatenate([q, v])
    a = arm.aba(q, v, tau)
    c = _eval(host_rb, 2, "c", model.params(), False, x, tau)
    np.testing.assert_allclose(c[5:7], a, rtol=1e-12, atol=1e-13)
    vn = v + dt * a
    np.testing.assert_allclose(c[1:5], np.concatenate([q + dt * vn, vn]), rtol=1e-13, atol=1e-14)
    d = _eval(host_rb, 2, "d", model.params(), False, x, tau)
    Fx = d[:16].reshape(4, 4, order="F")
    Fu = d[16:24].reshape(4, 2, order="F")
    wq, wv, wt = arm.aba_derivatives(q, v, tau)
    I2 = np.eye(2)
    want_Fx = np.block([[I2 + dt * dt * wq, dt * (I2 + dt * wv)], [dt * wq, I2 + dt * wv]])
    want_Fu = np.vstack([dt * dt * wt, dt * wt])
    np.testing.assert_allclose(Fx, want_Fx, rtol=1e-11, atol=1e-12)
    np.testing.assert_allclose(Fu, want_Fu, rtol=1e-11, atol=1e-13)


# ---------------------------------------------------------------- FP64 work per node on a counting scalar
def test_counting_scalar_flops_match_the_committed_record(tmp_path):
    """profiles/r02_arm_flops_counting_scalar.json ("ABA flops" per node) is what
    tests/flop_count/count_arm_flops.cpp prints for the synthetic 7-DoF arm; bench.py's arm roofline uses the smaller
    count of what the kernel issues (tools/count_fp64_ops.py), which must stay below this algorithmic count."""
    import json
    import os
    import subprocess
    here = os.path.dirname(os.path.abspath(__file__))
    exe = str(tmp_path / "count")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-I", os.path.join(here, "..", "oracle"), "-o", exe,
                           os.path.join(here, "flop_count", "count_arm_flops.cpp")])
    pfile = str(tmp_path / "params.bin")
    np.asarray(cb.ArmActionModel(cb.synthetic_arm(7)).params(), dtype=np.float64).tofile(pfile)
    got = json.loads(subprocess.check_output([exe, pfile, "7"], text=True))
    want = json.load(open(os.path.join(here, "..", "profiles", "r02_arm_flops_counting_scalar.json")))
    for key in ("calc", "calcdiff", "aba_alone", "aba_derivatives"):
        assert got[key] == want[key]
    import bench
    assert bench.Arm.flops_calc <= got["calc"]["flops"]
    assert bench.Arm.flops_calcdiff <= got["calcdiff"]["flops"]
