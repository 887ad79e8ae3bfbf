"""CPU self-consistency tests of the oracle's rigid-body arithmetic (oracle/rbd.hpp, arm.hpp). The reference holds
no golden values for Pinocchio's ABA / ABA derivatives ("parity unpinned"), so the restated algorithms are
pinned the way the reference's own tests pin this boundary: finite differences (unittest/test_diff_actions.cpp:103-135,
tol 3e4*sqrt(2 eps)), quasiStatic => zero acceleration (test_diff_actions.cpp:67-84), plus identities between
independently written algorithms (M qdd + b = tau with M from CRBA, b from RNEA; d qdd/d tau = M^-1 symmetric)."""
import numpy as np
import pytest
from scipy.linalg import expm

import crocoddyl_b200 as cb
import oracle_lib as ol

NV = 7
TOL_ND = 3e4 * np.sqrt(2 * np.finfo(float).eps)


def _random_chain(nv, seed):
    rng = np.random.default_rng(seed)
    axes = rng.standard_normal((nv, 3))
    axes /= np.linalg.norm(axes, axis=1)[:, None]
    R = []
    for _ in range(nv):
        Q, _r = np.linalg.qr(rng.standard_normal((3, 3)))
        if np.linalg.det(Q) < 0:
            Q[:, 0] *= -1
        R.append(Q)
    p = 0.3 * rng.standard_normal((nv, 3))
    I = []
    for _ in range(nv):
        A = rng.standard_normal((3, 3))
        I.append(0.01 * (A @ A.T + np.eye(3)))
    return cb.RevoluteChain(axes, R, p, rng.uniform(0.5, 3., nv), 0.1 * rng.standard_normal((nv, 3)), I)


def _model(kind, seed=0):
    chain = cb.synthetic_arm(NV) if kind == "synthetic" else _random_chain(NV, seed)
    return cb.ArmActionModel(chain, 1e-3)


def _rand_state(seed):
    rng = np.random.default_rng(seed)
    return rng.uniform(-1, 1, NV), rng.standard_normal(NV), rng.standard_normal(NV)


@pytest.mark.parametrize("kind", ["synthetic", "random"])
def test_aba_inverts_rnea_and_crba(kind):
    P = _model(kind).params()
    for s in range(4):
        q, v, tau = _rand_state(s)
        a = ol.arm_op(P, "aba", np.concatenate([q, v, tau]), NV)
        np.testing.assert_allclose(ol.arm_op(P, "rnea", np.concatenate([q, v, a]), NV), tau, rtol=1e-10, atol=1e-10)
        M = ol.arm_op(P, "crba", q, NV * NV).reshape(NV, NV, order="F")
        b = ol.arm_op(P, "rnea", np.concatenate([q, v, 0 * a]), NV)
        np.testing.assert_allclose(M, M.T, atol=1e-13)
        assert np.all(np.linalg.eigvalsh(M) > 0)
        np.testing.assert_allclose(M @ a + b, tau, rtol=1e-10, atol=1e-10)


@pytest.mark.parametrize("kind", ["synthetic", "random"])
def test_aba_derivatives_against_numdiff_and_minv(kind):
    P = _model(kind).params()
    q, v, tau = _rand_state(11)
    out = ol.arm_op(P, "aba_derivatives", np.concatenate([q, v, tau]), 3 * NV * NV)
    dq, dv, Minv = [out[i * NV * NV:(i + 1) * NV * NV].reshape(NV, NV, order="F") for i in range(3)]
    M = ol.arm_op(P, "crba", q, NV * NV).reshape(NV, NV, order="F")
    np.testing.assert_allclose(Minv, np.linalg.inv(M), rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(Minv, Minv.T, atol=1e-12)
    h = 1e-6
    f = lambda q_, v_, t_: ol.arm_op(P, "aba", np.concatenate([q_, v_, t_]), NV)
    for k in range(NV):
        e = np.zeros(NV)
        e[k] = h
        np.testing.assert_allclose((f(q + e, v, tau) - f(q - e, v, tau)) / (2 * h), dq[:, k], rtol=1e-6, atol=1e-6)
        np.testing.assert_allclose((f(q, v + e, tau) - f(q, v - e, tau)) / (2 * h), dv[:, k], rtol=1e-6, atol=1e-6)
        np.testing.assert_allclose((f(q, v, tau + e) - f(q, v, tau - e)) / (2 * h), Minv[:, k], rtol=1e-6, atol=1e-6)


def test_rnea_derivatives_against_numdiff():
    P = _model("random", 3).params()
    q, v, a = _rand_state(5)
    out = ol.arm_op(P, "rnea_derivatives", np.concatenate([q, v, a]), 2 * NV * NV)
    tq, tv = [out[i * NV * NV:(i + 1) * NV * NV].reshape(NV, NV, order="F") for i in range(2)]
    h = 1e-6
    f = lambda q_, v_: ol.arm_op(P, "rnea", np.concatenate([q_, v_, a]), NV)
    for k in range(NV):
        e = np.zeros(NV)
        e[k] = h
        np.testing.assert_allclose((f(q + e, v) - f(q - e, v)) / (2 * h), tq[:, k], rtol=1e-6, atol=1e-6)
        np.testing.assert_allclose((f(q, v + e) - f(q, v - e)) / (2 * h), tv[:, k], rtol=1e-6, atol=1e-6)


def test_pendulum_closed_form():
    """One link about Y, COM at (0,0,-l): qdd = -(m g l sin q) / (Iyy + m l^2) + tau / (Iyy + m l^2)."""
    m, l, Iyy, g = 1.7, 0.4, 0.03, 9.81
    chain = cb.RevoluteChain([(0., 1., 0.)], [np.eye(3)], [(0., 0., 0.)], [m], [(0., 0., -l)],
                             [np.diag([0.01, Iyy, 0.02])])
    P = cb.ArmActionModel(chain, 1e-3).params()
    for q, v, tau in [(0.3, 0.0, 0.0), (-1.1, 2.0, 0.5), (2.5, -1.0, -0.2)]:
        a = ol.arm_op(P, "aba", np.array([q, v, tau]), 1)[0]
        np.testing.assert_allclose(a, (tau - m * g * l * np.sin(q)) / (Iyy + m * l * l), rtol=1e-12)


def _se3(R, p):
    M = np.eye(4)
    M[:3, :3] = R
    M[:3, 3] = p
    return M


def _hat6(xi):  # [v; w] -> 4x4 twist
    v, w = xi[:3], xi[3:]
    W = np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]])
    T = np.zeros((4, 4))
    T[:3, :3] = W
    T[:3, 3] = v
    return T


def _log6(P, M):
    return ol.arm_op(P, "log6", np.concatenate([M[:3, :3].ravel(), M[:3, 3]]), 6)


def test_log6_inverts_exp6_and_jlog6_is_its_derivative():
    P = _model("synthetic").params()
    rng = np.random.default_rng(4)
    for scale in (1.0, 1e-3, 1e-6, 0.0, 2.5):
        xi = scale * rng.standard_normal(6)
        if np.linalg.norm(xi[3:]) > 3.0:
            xi[3:] *= 3.0 / np.linalg.norm(xi[3:])
        M = expm(_hat6(xi))
        np.testing.assert_allclose(_log6(P, M), xi, rtol=1e-9, atol=1e-12)
        J = ol.arm_op(P, "jlog6", np.concatenate([M[:3, :3].ravel(), M[:3, 3]]), 36).reshape(6, 6, order="F")
        h = 1e-6
        Jn = np.zeros((6, 6))
        for k in range(6):
            e = np.zeros(6)
            e[k] = h
            Jn[:, k] = (_log6(P, M @ expm(_hat6(e))) - _log6(P, M @ expm(_hat6(-e)))) / (2 * h)
        np.testing.assert_allclose(J, Jn, rtol=1e-6, atol=1e-6)
    # continuity across the Taylor threshold eps^(1/4)
    thr = np.finfo(float).eps ** 0.25
    ax = np.array([0.3, -0.5, 0.8]) / np.linalg.norm([0.3, -0.5, 0.8])
    for th in (thr * (1 - 1e-6), thr * (1 + 1e-6)):
        xi = np.concatenate([[0.2, -0.1, 0.4], th * ax])
        # acos((tr-1)/2) at theta ~ 1e-4 carries a relative error ~ eps/theta^2 (inherent to the published formula)
        np.testing.assert_allclose(_log6(P, expm(_hat6(xi))), xi, rtol=1e-6, atol=1e-12)


@pytest.mark.parametrize("kind", ["synthetic", "random"])
def test_frame_jacobian_local_against_numdiff(kind):
    mod = _model(kind)
    P = mod.params()
    q = _rand_state(9)[0]

    def placement(q_):
        o = ol.arm_op(P, "frame", q_, 12 + 6 * NV)
        return _se3(o[:9].reshape(3, 3), o[9:12]), o[12:].reshape(6, NV, order="F")

    M0, J = placement(q)
    h = 1e-6
    for k in range(NV):
        e = np.zeros(NV)
        e[k] = h
        Mp, _ = placement(q + e)
        Mm, _ = placement(q - e)
        col = (_log6(P, np.linalg.inv(M0) @ Mp) - _log6(P, np.linalg.inv(M0) @ Mm)) / (2 * h)
        np.testing.assert_allclose(J[:, k], col, rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("kind", ["synthetic", "random"])
@pytest.mark.parametrize("terminal", [False, True])
def test_arm_action_calcdiff_against_numdiff(kind, terminal):
    """FreeFwdDynamics + Euler calcDiff vs ActionModelNumDiff (unittest/test_diff_actions.cpp:103-135 pattern)."""
    mod = _model(kind, 2)
    P = mod.params()
    rng = np.random.default_rng(8)
    ob = ol.OracleBatch("arm", "fddp", 2, 2 * NV, NV, P, np.zeros((1, 2 * NV)))
    for _ in range(3):
        x = np.concatenate([rng.uniform(-1, 1, NV), rng.standard_normal(NV)])
        u = np.zeros(NV) if terminal else rng.standard_normal(NV)
        a, d = ob.numdiff(0, x, u, terminal=terminal)
        for k in ("Fx", "Lx") + (() if terminal else ("Fu", "Lu")):
            np.testing.assert_allclose(a[k], d[k], atol=TOL_ND, rtol=1e-4, err_msg=k)
        if terminal:
            np.testing.assert_allclose(a["Fx"], np.eye(2 * NV))
            np.testing.assert_allclose(a["Fu"], 0.)


def test_quasistatic_gives_zero_acceleration():
    mod = _model("random", 6)
    P = mod.params()
    q = _rand_state(1)[0]
    u = ol.arm_op(P, "quasistatic", q, NV)
    a = ol.arm_op(P, "aba", np.concatenate([q, np.zeros(NV), u]), NV)
    assert np.max(np.abs(a)) <= 1e-8  # test_diff_actions.cpp:67-84


def test_fddp_on_arm_converges_and_decreases_cost():
    """The reference's example trace (examples/log/arm_manipulation.log) converges in a handful of iterations from a
    quasi-static start; the oracle must show the same qualitative behaviour on the synthetic arm (config C4)."""
    mod = _model("synthetic")
    P = mod.params()
    T, B = 30, 3
    rng = np.random.default_rng(2)
    x0 = np.concatenate([rng.uniform(-1, 1, (B, NV)), np.zeros((B, NV))], 1)
    us0 = np.stack([ol.arm_op(P, "quasistatic", x0[b, :NV], NV) for b in range(B)])
    xs = np.repeat(x0[:, None, :], T + 1, 1)
    us = np.repeat(us0[:, None, :], T, 1)
    ob = ol.OracleBatch("arm", "fddp", T, 2 * NV, NV, P, x0)
    ob.solve(xs, us, 100, False, 0.1)
    for b in range(B):
        tr = ob.trace(b)
        assert tr.shape[0] >= 2
        assert tr[-1, 1] < tr[0, 1]
        assert np.all(np.isfinite(tr[:, 1]))
