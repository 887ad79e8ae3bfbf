"""ctypes binding of the C ABI (include/crocoddyl_b200.h -> crocoddyl_b200/lib/libcrocoddyl_b200.so).

There is no CPU fallback: if the CUDA library is missing this module raises, and every compute entry
point of the library itself fails without a CUDA device.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CROCODDYL_B200_LIB", os.path.join(HERE, "lib", "libcrocoddyl_b200.so"))

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_vp = C.c_void_p
_vpp = C.POINTER(C.c_void_p)
_cp = C.c_char_p
_up = C.POINTER(C.c_ubyte)


class IterRecord(C.Structure):
    _fields_ = [("problem", C.c_int), ("iter", C.c_int), ("cost", C.c_double), ("stop", C.c_double),
                ("d0", C.c_double), ("d1", C.c_double), ("xreg", C.c_double), ("ureg", C.c_double),
                ("steplength", C.c_double), ("dV", C.c_double), ("dVexp", C.c_double), ("is_feasible", C.c_int),
                ("alpha_index", C.c_int), ("retries", C.c_int)]


CALLBACK = C.CFUNCTYPE(None, C.c_void_p, C.POINTER(IterRecord))

# every symbol include/crocoddyl_b200.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "cgx_last_error": (_cp, []),
    "cgx_device_count": (C.c_int, []),
    "cgx_set_device": (C.c_int, [C.c_int]),
    "cgx_version": (_cp, []),
    "cgx_problem_create": (C.c_int, [_vpp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _dp, C.c_long, C.c_int, _dp]),
    "cgx_problem_destroy": (C.c_int, [_vp]),
    "cgx_problem_set_terminal_params": (C.c_int, [_vp, _dp, C.c_long]),
    "cgx_problem_get_terminal_params": (C.c_int, [_vp, _dp]),
    "cgx_problem_set_u_bounds": (C.c_int, [_vp, _dp, _dp, C.c_int]),
    "cgx_problem_get_u_bounds": (C.c_int, [_vp, _dp, _dp, _ip, _ip]),
    "cgx_problem_set_x0": (C.c_int, [_vp, _dp]),
    "cgx_problem_get_x0": (C.c_int, [_vp, _dp]),
    "cgx_problem_dims": (C.c_int, [_vp, _ip, _ip, _ip, _ip, _ip]),
    "cgx_problem_calc": (C.c_int, [_vp, _dp, _dp, _dp]),
    "cgx_problem_calc_diff": (C.c_int, [_vp, _dp, _dp, _dp]),
    "cgx_problem_rollout": (C.c_int, [_vp, _dp, _dp]),
    "cgx_problem_quasi_static": (C.c_int, [_vp, _dp, _dp]),
    "cgx_problem_get_data": (C.c_int, [_vp, _cp, C.c_int, C.c_int, _dp]),
    "cgx_arm_param_count": (C.c_long, [C.c_int]),
    "cgx_solver_create": (C.c_int, [_vpp, _vp, C.c_int]),
    "cgx_solver_destroy": (C.c_int, [_vp]),
    "cgx_solver_set_param": (C.c_int, [_vp, _cp, C.c_double]),
    "cgx_solver_get_param": (C.c_int, [_vp, _cp, _dp]),
    "cgx_solver_set_alphas": (C.c_int, [_vp, _dp, C.c_int]),
    "cgx_solver_get_alphas": (C.c_int, [_vp, _dp, _ip]),
    "cgx_solver_set_storage": (C.c_int, [_vp, C.c_int]),
    "cgx_solver_enable_trace": (C.c_int, [_vp, C.c_int, C.c_int]),
    "cgx_solver_solve": (C.c_int, [_vp, _dp, _dp, C.c_int, C.c_int, C.c_double, _up]),
    "cgx_solver_solve_fetch": (C.c_int, [_vp, _dp, _dp, C.c_int, C.c_int, C.c_double, _up, _dp, _dp]),
    "cgx_solver_upload": (C.c_int, [_vp, _dp, _dp]),
    "cgx_solver_solve_resident": (C.c_int, [_vp, C.c_int, C.c_int, C.c_double]),
    "cgx_solver_download": (C.c_int, [_vp, _up]),
    "cgx_solver_set_candidate": (C.c_int, [_vp, _dp, _dp, C.c_int]),
    "cgx_solver_compute_direction": (C.c_int, [_vp, C.c_int, _ip]),
    "cgx_solver_try_step": (C.c_int, [_vp, C.c_double, _dp, _ip]),
    "cgx_solver_expected_improvement": (C.c_int, [_vp, _dp]),
    "cgx_solver_stopping_criteria": (C.c_int, [_vp, _dp]),
    "cgx_solver_get": (C.c_int, [_vp, _cp, C.c_int, C.c_int, _dp]),
    "cgx_solver_get_all": (C.c_int, [_vp, _cp, _dp]),
    "cgx_solver_set_xs": (C.c_int, [_vp, _dp]),
    "cgx_solver_set_us": (C.c_int, [_vp, _dp]),
    "cgx_solver_get_scalars": (C.c_int, [_vp, _dp]),
    "cgx_solver_get_scalars_row": (C.c_int, [_vp, C.c_int, _dp]),
    "cgx_solver_get_trace": (C.c_int, [_vp, C.c_int, _dp, C.c_int]),
    "cgx_solver_set_callback": (C.c_int, [_vp, CALLBACK, _vp]),
    "cgx_solver_last_timing": (C.c_int, [_vp, _dp, _ip]),
    "cgx_probe_fp64_tflops": (C.c_int, [_dp]),
    "cgx_unicycle_calc_noalloc": (C.c_int, [C.c_int, C.c_int, _dp, _dp, _dp, _dp, _dp]),
}

NOALLOC_SCRATCH_DOUBLES = 1048576  # CGX_NOALLOC_SCRATCH_DOUBLES

NSCALARS = 17
NTRACE = 13
SCALARS = ["cost", "stop", "d0", "d1", "xreg", "ureg", "steplength", "dV", "dVexp", "iter", "is_feasible",
           "converged", "n_records", "cost_try", "was_feasible", "n_backward", "n_forward"]
TRACE = ["iter", "cost", "stop", "d0", "d1", "xreg", "ureg", "steplength", "is_feasible", "dV", "dVexp",
         "alpha_idx", "retries"]

_lib = None


class CgxError(RuntimeError):
    """Raised for every non-zero status of the C ABI (text = cgx_last_error(), which mirrors the
    reference's crocoddyl::Exception messages)."""

    def __init__(self, code, msg):
        RuntimeError.__init__(self, msg)
        self.code = code


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("crocoddyl_b200: %s is missing; run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(there is no CPU fallback)" % LIB_PATH)
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(rc):
    if rc < 0:
        raise CgxError(rc, lib().cgx_last_error().decode())
    return rc


def dptr(a):
    return None if a is None else a.ctypes.data_as(_dp)
