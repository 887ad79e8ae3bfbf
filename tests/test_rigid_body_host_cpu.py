"""The product's rigid-body arithmetic (crocoddyl_b200/csrc/rigid_body.cuh, __host__ __device__) instantiated on the
host (tests/host_rb/host_rb.cpp, g++) and compared with the oracle's independently written statements
(oracle/rbd.hpp: 6x6 articulated inertias, CRBA + Cholesky, dense Jlog6 product) on the same inputs. This is what can
be checked of the arm kernels without a GPU; tests/test_arm_gpu.py repeats it through the C ABI on the device."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import crocoddyl_b200 as cb
import oracle_lib as ol
from test_oracle_arm_cpu import _random_chain

HERE = os.path.dirname(os.path.abspath(__file__))
_dp = C.POINTER(C.c_double)


@pytest.fixture(scope="module")
def host_rb():
    out_dir = os.path.join(HERE, "host_rb", "_build")
    os.makedirs(out_dir, exist_ok=True)
    so = os.path.join(out_dir, "libhost_rb.so")
    src = os.path.join(HERE, "host_rb", "host_rb.cpp")
    hdr = os.path.join(HERE, "..", "crocoddyl_b200", "csrc", "rigid_body.cuh")
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-o", so, src])
    lib = C.CDLL(so)
    lib.host_rb_eval.argtypes = [C.c_int, C.c_char, _dp, C.c_int, _dp, _dp, _dp]
    return lib


def _eval(lib, nv, op, params, terminal, x, u):
    out = np.zeros(4096)
    p = np.ascontiguousarray(params, dtype=np.float64)
    x = np.ascontiguousarray(x, dtype=np.float64)
    up = None if u is None else np.ascontiguousarray(u, dtype=np.float64).ctypes.data_as(_dp)
    n = lib.host_rb_eval(nv, op.encode(), p.ctypes.data_as(_dp), int(terminal), x.ctypes.data_as(_dp), up,
                         out.ctypes.data_as(_dp))
    assert n > 0
    return out[:n]


def _model(kind, nv, seed=0):
    chain = cb.synthetic_arm(nv) if kind == "synthetic" else _random_chain(nv, seed)
    return cb.ArmActionModel(chain, 1e-3, w_gripperPose=1., w_xReg=1e-2, w_uReg=1e-3)


@pytest.mark.parametrize("kind,nv", [("synthetic", 7), ("random", 7), ("random", 4), ("random", 2), ("synthetic", 1)])
def test_node_calc_and_calcdiff_match_oracle(host_rb, kind, nv):
    model = _model(kind, nv, seed=3)
    params = model.params()
    n = 2 * nv
    T = 3
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
        # the terminal node is evaluated on its own parameter block (the product's default: the running block with dt = 0)
        pt = model.terminal_model.params() if term else params
        got = _eval(host_rb, nv, "c", pt, term, xs[t], u)
        np.testing.assert_allclose(got[0], ob.get("node_cost", 0, t)[0], rtol=1e-12, atol=1e-14)
        np.testing.assert_allclose(got[1:1 + n], ob.get("xnext", 0, t), rtol=1e-12, atol=1e-13)
        d = _eval(host_rb, nv, "d", pt, term, xs[t], u)
        o = 0
        for name, shp in [("Fx", (n, n)), ("Fu", (n, nv)), ("Lx", (n,)), ("Lu", (nv,)), ("Lxx", (n, n)), ("Luu", (nv, nv))]:
            sz = int(np.prod(shp))
            val = d[o:o + sz].reshape(shp, order="F")
            o += sz
            if term and name in ("Fu", "Lu", "Luu"):
                if name == "Fu":
                    assert np.all(val == 0.)
                continue
            ref = ob.get(name, 0, t)
            np.testing.assert_allclose(val, ref, rtol=1e-10, atol=1e-11 * max(1., np.abs(ref).max()), err_msg=name)


@pytest.mark.parametrize("kind,nv", [("synthetic", 7), ("random", 6), ("random", 3)])
def test_quasi_static_and_qdd_match_oracle(host_rb, kind, nv):
    model = _model(kind, nv, seed=5)
    params = model.params()
    rng = np.random.default_rng(2)
    for _ in range(5):
        q, v, tau = rng.uniform(-2, 2, nv), rng.standard_normal(nv), 5. * rng.standard_normal(nv)
        x = np.concatenate([q, v])
        uq = _eval(host_rb, nv, "q", params, False, x, None)
        np.testing.assert_allclose(uq, ol.arm_op(params, "quasistatic", q, nv), rtol=1e-12, atol=1e-13)
        got = _eval(host_rb, nv, "c", params, False, x, tau)
        qdd = got[1 + 2 * nv:]
        ref = ol.arm_op(params, "aba", np.concatenate([q, v, tau]), nv)
        np.testing.assert_allclose(qdd, ref, rtol=1e-11, atol=1e-11 * max(1., np.abs(ref).max()))


def test_frame_cost_near_zero_rotation_and_near_pi(host_rb):
    """log6 / Jlog6 branches: Mref chosen so that rMf is (almost) the identity and (almost) a half turn."""
    nv = 2
    chain = cb.synthetic_arm(nv)
    x = np.array([0., 0., 0.3, -0.2])
    for ang in (0., 1e-6, np.pi - 1e-3, 3.0):
        c, s = np.cos(ang), np.sin(ang)
        Rref = np.array([[c, -s, 0.], [s, c, 0.], [0., 0., 1.]])
        model = cb.ArmActionModel(chain, 1e-3, Mref_R=Rref, Mref_p=(0.1, 0., 0.5))
        params = model.params()
        ob = ol.OracleBatch("arm", "fddp", 1, 2 * nv, nv, params, x[None])
        ob.set_candidate(np.stack([x, x])[None], np.zeros((1, 1, nv)), False)
        ob.calc()
        ob.calcdiff()
        got = _eval(host_rb, nv, "c", params, False, x, np.zeros(nv))
        np.testing.assert_allclose(got[0], ob.get("node_cost", 0, 0)[0], rtol=1e-11)
        d = _eval(host_rb, nv, "d", params, False, x, np.zeros(nv))
        n = 2 * nv
        Lx = d[n * n + n * nv:n * n + n * nv + n]
        np.testing.assert_allclose(Lx, ob.get("Lx", 0, 0), rtol=1e-9, atol=1e-12)
