"""CPU checks of the C-ABI library: it loads, exports every symbol include/crocoddyl_b200.h declares,
and (no CPU fallback) every compute entry point fails loudly without a CUDA device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import crocoddyl_b200 as cb
from crocoddyl_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "crocoddyl_b200.h")).read()
    return re.findall(r"^CGX_API [a-z* ]+ \*?(cgx_[a-z0-9_]+)\(", src, flags=re.M) + ["cgx_last_error"]


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(_lib.LIB_PATH)
    names = _header_symbols()
    assert len(names) >= 40
    for n in names:
        assert hasattr(lib, n), n
    assert set(names) == set(_lib.SIGNATURES), set(names) ^ set(_lib.SIGNATURES)


def test_header_cites_reference_for_entry_points():
    src = open(os.path.join(ROOT, "include", "crocoddyl_b200.h")).read()
    for ref in ("ddp.cpp:41-118", "fddp.cpp:19-105", "shooting.hxx:133-161", "solver-base.cpp:42-67"):
        assert ref in src


def test_no_cpu_fallback():
    if cb.device_count() > 0:
        pytest.skip("a CUDA device is present")
    m = cb.ActionModelUnicycle()
    with pytest.raises(cb.CgxError) as e:
        cb.ShootingProblem(np.zeros(3), [m] * 5, m)
    assert e.value.code == -2
    assert "no CPU fallback" in str(e.value)


def test_argument_validation_mirrors_reference_messages():
    m = cb.ActionModelUnicycle()
    with pytest.raises(cb.CgxError) as e:
        cb.ShootingProblem(np.zeros(4), [m] * 5, m)
    assert "x0 has wrong dimension" in str(e.value)
    m2 = cb.ActionModelUnicycle()
    with pytest.raises(cb.CgxError):
        cb.ShootingProblem(np.zeros(3), [m, m2], m)


def test_lqr_param_packing_is_column_major():
    m = cb.ActionModelLQR(3, 2)
    m.Fu = np.arange(6.).reshape(3, 2)
    p = m.params()
    assert p.shape == (1 + 9 + 6 + 3 + 9 + 6 + 4 + 3 + 2,)
    np.testing.assert_array_equal(p[10:16], m.Fu.ravel(order="F"))


def test_null_solver_handle_is_an_invalid_argument_not_a_crash():
    """solve / solve_fetch on a NULL handle return CGX_ERR_INVALID with a message (no exception crosses the C ABI)."""
    L = cb.lib()
    conv = np.zeros(1, dtype=np.uint8)
    up = conv.ctypes.data_as(_lib._up)
    assert L.cgx_solver_solve(None, None, None, 1, 0, 1e-9, up) == -1
    assert L.cgx_solver_solve_fetch(None, None, None, 1, 0, 1e-9, up, None, None) == -1
    assert b"NULL" in L.cgx_last_error()
