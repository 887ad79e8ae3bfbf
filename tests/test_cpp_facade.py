"""C++ facade (include/crocoddyl_b200.hpp) over the C ABI: compiles with g++, refuses to run without a GPU (no CPU
fallback), and on a GPU reproduces the oracle's FDDP solve of the benchmark/unicycle-optctrl.cpp problem."""
import os
import subprocess

import numpy as np
import pytest

import crocoddyl_b200 as cb
import oracle_lib as ol

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build():
    out_dir = os.path.join(ROOT, "tests", "cpp", "_build")
    os.makedirs(out_dir, exist_ok=True)
    exe = os.path.join(out_dir, "facade_demo")
    src = os.path.join(ROOT, "tests", "cpp", "facade_demo.cpp")
    lib = cb._lib.LIB_PATH
    cb.lib()  # builds / loads the library
    deps = [src, os.path.join(ROOT, "include", "crocoddyl_b200.hpp"), os.path.join(ROOT, "include", "crocoddyl_b200.h")]
    if not os.path.exists(exe) or os.path.getmtime(exe) < max(os.path.getmtime(d) for d in deps):
        subprocess.check_call(["g++", "-std=c++17", "-O1", "-I", os.path.join(ROOT, "include"), src, "-o", exe,
                               "-L", os.path.dirname(lib), "-lcrocoddyl_b200", "-Wl,-rpath," + os.path.dirname(lib)])
    return exe


def test_facade_compiles_and_refuses_without_gpu():
    exe = _build()
    if cb.device_count() > 0:
        pytest.skip("a GPU is present: covered by the gpu test")
    r = subprocess.run([exe, "10"], stdout=subprocess.PIPE, text=True)
    assert r.returncode == 3, r.stdout
    assert "no CUDA device" in r.stdout or "CUDA" in r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("kind,T", [("unicycle", 20), ("lqr", 15), ("euler_difflqr", 12)])
def test_facade_solve_matches_oracle(kind, T):
    exe = _build()
    arg = {"unicycle": [], "lqr": ["lqr"], "euler_difflqr": ["difflqr"]}[kind]
    r = subprocess.run([exe, str(T)] + arg, stdout=subprocess.PIPE, text=True)
    assert r.returncode == 0, r.stdout
    lines = r.stdout.strip().split("\n")
    head = lines[0].split()
    vals = {ln.split()[0]: np.array([float(v) for v in ln.split()[1:]]) for ln in lines[1:4]}
    if kind == "unicycle":
        nx, nu, params, x0 = 3, 2, ol.unicycle_params(), np.array([[-1., -1., 1.]])
    elif kind == "lqr":
        nx, nu, params, x0 = 6, 2, ol.lqr_params(6, 2), np.zeros((1, 6))
    else:  # the oracle evaluates the differential model first and integrates after, as the reference does
        from integrated_cases import oracle_params
        d = cb.DifferentialActionModelLQR(3, 2, False)
        for name in ("Fq", "Fv", "Fu", "f0", "Lxx", "Lxu", "Luu", "lx", "lu"):
            setattr(d, name, np.asarray(getattr(d, name))[None])
        nx, nu, params, x0 = 6, 2, oracle_params(d, 0.05)[0], np.zeros((1, 6))
    ob = ol.OracleBatch(kind, "fddp", T, nx, nu, params, x0)
    ob.solve(np.repeat(x0[:, None, :], T + 1, 1), np.zeros((1, T, nu)), 100, False, 1e-9)
    so = ob.scalars()
    assert int(head[1]) == int(so["converged"][0]) and int(head[3]) == int(so["iter"][0])
    np.testing.assert_allclose(float(head[5]), so["cost"][0], rtol=1e-9)
    np.testing.assert_allclose(vals["xs"], ob.get_all("xs").ravel(), rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(vals["us"], ob.get_all("us").ravel(), rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(vals["K0"], ob.get("K", 0, 0).ravel(order="F"), rtol=1e-9, atol=1e-9)
    assert "Warm start state has wrong dimension" in lines[4]
    # node plugin interface: model->calc / calcDiff on one node against the oracle's node model
    x = 0.3 + 0.1 * np.arange(nx)
    u = -0.2 + 0.5 * np.arange(nu)
    one = ol.OracleBatch(kind, "ddp", 1, nx, nu, params, x[None])
    one.set_candidate(np.stack([x, x])[None], u[None, None], False)
    one.calc()
    one.calcdiff()
    node = {ln.split()[0]: np.array([float(v) for v in ln.split()[1:]]) for ln in lines[6:9]}
    np.testing.assert_allclose(float(lines[5].split()[2]), one.get("node_cost", 0, 0)[0], rtol=1e-12)
    np.testing.assert_allclose(node["node_xnext"], one.get("xnext", 0, 0), rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(node["node_Fx"], one.get("Fx", 0, 0).ravel(order="F"), rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(node["node_Lu"], one.get("Lu", 0, 0), rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(float(lines[9].split()[2]), one.get("node_cost", 0, 1)[0], rtol=1e-12)
    # control-limited solve through SolverBoxFDDP
    bo = ol.OracleBatch(kind, "boxfddp", T, nx, nu, params, x0)
    bo.set_u_bounds(-0.4 * np.ones(nu), 0.4 * np.ones(nu))
    bo.solve(np.repeat(x0[:, None, :], T + 1, 1), np.zeros((1, T, nu)), 100, False, 1e-9)
    bh = lines[10].split()
    assert bh[0] == "box" and float(bh[8]) == 5e-5
    assert int(bh[2]) == int(bo.scalars()["converged"][0]) and int(bh[4]) == int(bo.scalars()["iter"][0])
    np.testing.assert_allclose(float(bh[6]), bo.scalars()["cost"][0], rtol=1e-9)
    bu = np.array([float(v) for v in lines[11].split()[1:]])
    assert np.all(np.abs(bu) <= 0.4 + 1e-15)
    np.testing.assert_allclose(bu, bo.get_all("us").ravel(), rtol=1e-7, atol=1e-8)
