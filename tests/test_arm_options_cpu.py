"""SURVEY 8 f3 on the arm model: the armature branch of DifferentialActionModelFreeFwdDynamics
(free-fwddyn.hxx:66-74,108-114), ActivationModelWeightedQuad on the state cost (weighted-quadratic.hpp:39-63) and
CostModelFrameTranslation / CostModelFrameRotation (frame-translation.hxx:51-79, frame-rotation.hxx:51-85) as the goal
cost. CPU side: the oracle's statements against closed forms / finite differences (the reference's own test structure,
unittest/test_diff_actions.cpp:103-135, tolerance 3e4 sqrt(2 eps)), and the product's rigid_body.cuh instantiated on the
host against the oracle. The oracle solves the armature case the way the reference does (CRBA + armature -> Minv ->
Minv (u - nle)); the product adds the armature to the joint pivots of its articulated-body sweep — two independent
routes to the same acceleration."""
import numpy as np
import pytest

import crocoddyl_b200 as cb
import oracle_lib as ol
from test_oracle_arm_cpu import _random_chain
from test_rigid_body_host_cpu import _eval, host_rb  # noqa: F401  (fixture)

TOL_ND = 3e4 * np.sqrt(2 * np.finfo(float).eps)

OPTIONS = {
    "armature": dict(armature=[0.1, 0.2, 0.05, 0.3, 0.15, 0.02, 0.07]),
    "weighted": dict(xWeights=np.concatenate([np.linspace(0.5, 3., 7), np.linspace(2., 0.1, 7)]), w_xReg=0.3),
    "translation": dict(goal="translation"),
    "rotation": dict(goal="rotation"),
    "all": dict(armature=[0.1] * 7, xWeights=np.linspace(0.2, 2., 14), w_xReg=0.1, goal="translation"),
}


def _model(name, nv=7, dt=1e-2, seed=1):
    kw = dict(OPTIONS[name])
    if nv != 7:
        if "armature" in kw:
            kw["armature"] = kw["armature"][:nv]
        if "xWeights" in kw:
            kw["xWeights"] = np.concatenate([np.asarray(kw["xWeights"])[:nv], np.asarray(kw["xWeights"])[7:7 + nv]])
    rng = np.random.default_rng(seed)
    Q, _r = np.linalg.qr(rng.standard_normal((3, 3)))
    if np.linalg.det(Q) < 0:
        Q[:, 0] *= -1
    kw.setdefault("w_xReg", 1e-2)
    return cb.ArmActionModel(_random_chain(nv, seed), dt, Mref_R=Q, Mref_p=(0.2, -0.1, 0.5), w_gripperPose=2., w_uReg=1e-2, **kw)


def test_pendulum_with_armature_closed_form(host_rb):  # noqa: F811
    """qdd = (tau - m g l sin q) / (Iyy + m l^2 + armature), oracle (Minv route) and product (ABA route)."""
    m, l, Iyy, g, arm = 1.7, 0.4, 0.03, 9.81, 0.25
    chain = cb.RevoluteChain([(0., 1., 0.)], [np.eye(3)], [(0., 0., 0.)], [m], [(0., 0., -l)], [np.diag([0.01, Iyy, 0.02])])
    model = cb.ArmActionModel(chain, 1e-3, armature=[arm])
    for q, v, tau in [(0.3, 0.0, 0.0), (-1.1, 2.0, 0.5), (2.5, -1.0, -0.2)]:
        want = (tau - m * g * l * np.sin(q)) / (Iyy + m * l * l + arm)
        x = np.array([[q, v]])
        ob = ol.OracleBatch("arm", "fddp", 1, 2, 1, model.params(), x)
        ob.set_candidate(np.stack([x[0], x[0]])[None], np.array([[[tau]]]), False)
        ob.calc()
        xn = ob.get("xnext", 0, 0)
        np.testing.assert_allclose((xn[1] - v) / 1e-3, want, rtol=1e-9)
        got = _eval(host_rb, 1, "c", model.params(), False, x[0], np.array([tau]))
        np.testing.assert_allclose(got[3], want, rtol=1e-12)


@pytest.mark.parametrize("name", sorted(OPTIONS))
@pytest.mark.parametrize("terminal", [False, True])
def test_oracle_calcdiff_against_numdiff(name, terminal):
    model = _model(name)
    nv, n = 7, 14
    P = (model.terminal_model if terminal else model).params()
    rng = np.random.default_rng(3)
    x = np.concatenate([rng.uniform(-1, 1, nv), rng.standard_normal(nv)])
    u = rng.standard_normal(nv)

    def node(x_, u_):
        ob = ol.OracleBatch("arm", "fddp", 1, n, nv, model.params(), x_[None], tparams=P)
        ob.set_candidate(np.stack([x_, x_])[None], u_[None, None], False)
        ob.calc()
        ob.calcdiff()
        return ob

    t = 1 if terminal else 0
    ob = node(x, u)
    Fx, Fu, Lx, Lu = ob.get("Fx", 0, t), ob.get("Fu", 0, t), ob.get("Lx", 0, t), ob.get("Lu", 0, t)
    h = 1e-6
    f = lambda x_, u_: (node(x_, u_).get("xnext", 0, t), node(x_, u_).get("node_cost", 0, t)[0])  # noqa: E731
    for k in range(n):
        e = np.zeros(n)
        e[k] = h
        (xp, cp), (xm, cm) = f(x + e, u), f(x - e, u)
        np.testing.assert_allclose((xp - xm) / (2 * h), Fx[:, k], rtol=TOL_ND, atol=TOL_ND)
        np.testing.assert_allclose((cp - cm) / (2 * h), Lx[k], rtol=TOL_ND, atol=TOL_ND)
    if not terminal:
        for k in range(nv):
            e = np.zeros(nv)
            e[k] = h
            (xp, cp), (xm, cm) = f(x, u + e), f(x, u - e)
            np.testing.assert_allclose((xp - xm) / (2 * h), Fu[:, k], rtol=TOL_ND, atol=TOL_ND)
            np.testing.assert_allclose((cp - cm) / (2 * h), Lu[k], rtol=TOL_ND, atol=TOL_ND)
    # Gauss-Newton Hessian of the goal cost + the state weights: Lxx is what the residual Jacobians give
    Lxx = ob.get("Lxx", 0, t)
    np.testing.assert_allclose(Lxx, Lxx.T, atol=1e-12)
    assert np.all(np.linalg.eigvalsh(Lxx) > -1e-12)


@pytest.mark.parametrize("name", sorted(OPTIONS))
@pytest.mark.parametrize("nv", [7, 2])
def test_product_node_matches_oracle(host_rb, name, nv):  # noqa: F811
    model = _model(name, nv)
    params, n, T = model.params(), 2 * nv, 2
    rng = np.random.default_rng(11)
    xs = np.concatenate([rng.uniform(-1.5, 1.5, (T + 1, nv)), rng.standard_normal((T + 1, nv))], 1)
    us = 3. * rng.standard_normal((T, nv))
    ob = ol.OracleBatch("arm", "fddp", T, n, nv, params, xs[:1])
    ob.set_candidate(xs[None], us[None], False)
    ob.calc()
    ob.calcdiff()
    for t in range(T + 1):
        term = t == T
        u = None if term else us[t]
        pt = model.terminal_model.params() if term else params
        got = _eval(host_rb, nv, "c", pt, term, xs[t], u)
        np.testing.assert_allclose(got[0], ob.get("node_cost", 0, t)[0], rtol=1e-11, atol=1e-13)
        np.testing.assert_allclose(got[1:1 + n], ob.get("xnext", 0, t), rtol=1e-11, atol=1e-12)
        d = _eval(host_rb, nv, "d", pt, term, xs[t], u)
        o = 0
        for fname, shp in [("Fx", (n, n)), ("Fu", (n, nv)), ("Lx", (n,)), ("Lu", (nv,)), ("Lxx", (n, n)), ("Luu", (nv, nv))]:
            sz = int(np.prod(shp))
            val = d[o:o + sz].reshape(shp, order="F")
            o += sz
            if term and fname in ("Fu", "Lu", "Luu"):
                continue
            want = ob.get(fname, 0, t)
            np.testing.assert_allclose(val, want, rtol=1e-9, atol=1e-10 * max(1., np.abs(want).max()), err_msg=fname)


def test_translation_goal_cost_is_half_squared_distance(host_rb):  # noqa: F811
    """CostModelFrameTranslation: 0.5 |p_frame(q) - p_ref|^2 with forward kinematics by plain 4x4 products."""
    from test_arm_thirdparty_cpu import _fk
    model = _model("translation")
    model = cb.ArmActionModel(model.chain, 1e-2, Mref_p=(0.2, -0.1, 0.5), w_gripperPose=1., w_xReg=0., w_uReg=0., goal="translation")
    rng = np.random.default_rng(2)
    for _ in range(3):
        q = rng.uniform(-2, 2, 7)
        x = np.concatenate([q, rng.standard_normal(7)])
        want = 0.5 * np.sum((_fk(model, q)[:3, 3] - model.Mref_p) ** 2)
        got = _eval(host_rb, 7, "c", model.terminal_model.params(), True, x, None)[0]
        np.testing.assert_allclose(got, want, rtol=1e-12)
