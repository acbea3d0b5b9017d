"""ctypes binding of libmpcb200.so (C ABI declared in include/mpcb200.h).

This is the only caller of the shared library.  There is no CPU fallback: if the
library has not been built, or the machine has no CUDA device, the product path raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# MPCB_LIB: another build of the same ABI (A/B timing of kernel variants); default is the in-tree library
LIB_PATH = os.environ.get("MPCB_LIB") or os.path.join(_HERE, "libmpcb200.so")

# model / road / flag constants (include/mpcb200.h)
MODEL_PACEJKA6, MODEL_KINEMATIC4 = 0, 1
ROAD_SHARED, ROAD_TABLE, ROAD_PER_INSTANCE, ROAD_BEZIER = 0, 1, 2, 3
FLAG_LANE, FLAG_OBST, FLAG_BOUNDED_D, FLAG_EXACT, FLAG_SAFE_DIST = 1, 2, 4, 8, 16
STATUS_UNKNOWN, STATUS_CONVERGED, STATUS_MAX_TIME, STATUS_MAX_ITER = 0, 1, 2, 3
STATUS_NOT_FINITE, STATUS_NO_PROGRESS, STATUS_INTERRUPTED = 4, 5, 6
E_ARG, E_CUDA, E_UNSUPPORTED, E_NODEVICE = -1, -2, -3, -4
NP, OBS_DIM, MAX_N, MAX_OBS = 22, 5, 64, 8
STATIC_SCRATCH_BYTES, STATIC_MAX_B = 32 << 20, 1 << 20  # mpcb_value_grad_static


class MpcProblemDesc(C.Structure):
    _fields_ = [
        ("model_id", C.c_int32), ("N", C.c_int32), ("S", C.c_int32), ("n_sub", C.c_int32),
        ("road_mode", C.c_int32), ("n_roads", C.c_int32), ("flags", C.c_int32), ("n_obs", C.c_int32),
        ("Ts", C.c_double), ("v_ref", C.c_double), ("c", C.c_double * 6),
        ("kin_friction", C.c_double), ("kin_accel", C.c_double),
        ("lane_a", C.c_double), ("lane_y_t", C.c_double),
        ("lane_b", C.c_double), ("lane_y_br", C.c_double), ("lane_y_bl", C.c_double),
        ("obs_Af", C.c_double), ("obs_b", C.c_double), ("obs_sigma_x", C.c_double),
        ("obs_sigma_y", C.c_double), ("obs_alpha_div", C.c_double),
        ("g_offset", C.c_double * 6), ("D_lower", C.c_double * 6), ("D_upper", C.c_double * 6),
        ("safe_x0", C.c_double), ("safe_y0", C.c_double), ("safe_ax_max", C.c_double), ("safe_ay_max", C.c_double),
    ]


class MpcSolverParams(C.Structure):
    _fields_ = [
        ("max_iter", C.c_int32), ("lbfgs_mem", C.c_int32), ("max_no_progress", C.c_int32),
        ("hess_vec_fd", C.c_int32),
        ("Lgamma_factor", C.c_double), ("L_min", C.c_double), ("L_max", C.c_double),
        ("lip_eps", C.c_double), ("lip_delta", C.c_double), ("tau_min", C.c_double),
        ("quad_tol", C.c_double), ("ls_tol", C.c_double),
        ("alm_max_iter", C.c_int32), ("reserved0", C.c_int32),
        ("eps_target", C.c_double), ("delta_target", C.c_double), ("Sigma0", C.c_double),
        ("eps0", C.c_double), ("rho", C.c_double), ("theta", C.c_double), ("Delta", C.c_double),
        ("M", C.c_double), ("Sigma_max", C.c_double),
    ]


_P = C.c_void_p  # device or host pointer, passed as an integer address
_I64 = C.c_int64

# name -> (restype, argtypes); every symbol include/mpcb200.h declares
SIGNATURES = {
    "mpcb_version": (C.c_int, []),
    "mpcb_default_desc": (None, [C.POINTER(MpcProblemDesc), C.c_int32, C.c_int32, C.c_int32]),
    "mpcb_default_solver_params": (None, [C.POINTER(MpcSolverParams), C.c_int32]),
    "mpcb_default_vehicle_params": (None, [C.POINTER(C.c_double)]),
    "mpcb_debug_timers": (C.c_int64, [C.c_void_p, C.POINTER(C.c_longlong), C.c_int64]),
    "mpcb_create": (C.c_int, [C.POINTER(MpcProblemDesc), C.POINTER(C.c_void_p)]),
    "mpcb_destroy": (None, [C.c_void_p]),
    "mpcb_last_error": (C.c_char_p, [C.c_void_p]),
    "mpcb_dynamics": (C.c_int, [C.c_void_p, _I64, _P, _P, _P, C.c_int32, _P, _P]),
    "mpcb_road_errors": (C.c_int, [C.c_void_p, _I64, _P, _P, _P, _P, C.c_int32, _P, _P, _P]),
    "mpcb_bezier_sample": (C.c_int, [C.c_void_p, _I64, _P, _P, _P]),
    "mpcb_rollout": (C.c_int, [C.c_void_p, _I64, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "mpcb_value_grad": (C.c_int, [C.c_void_p, _I64, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "mpcb_value_grad_static": (C.c_int, [C.c_void_p, _I64, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "mpcb_solve": (C.c_int, [C.c_void_p, C.POINTER(MpcSolverParams), _I64] + [_P] * 15),
    "mpcb_solve_host": (C.c_int, [C.c_void_p, C.POINTER(MpcSolverParams), _I64] + [_P] * 14),
    "mpcb_default_car_params": (None, [C.POINTER(C.c_double)]),
    "mpcb_game_payoff": (C.c_int, [C.c_void_p, _I64, C.c_int32, _P, _P, _P, _P, _P, _P, _P, _P]),
    "mpcb_check_guards": (_I64, [C.c_void_p]),
    "mpcb_last_finish_ns": (C.c_int, [C.c_void_p, _I64, _P, _P]),
    "mpcb_set_trace": (C.c_int, [C.c_void_p, C.c_int32]),
    "mpcb_last_trace": (C.c_int, [C.c_void_p, _I64, _P, _P]),
    "mpcb_last_inner_failures": (C.c_int, [C.c_void_p, _I64, _P, _P]),
    "mpcb_last_final_gamma": (C.c_int, [C.c_void_p, _I64, _P, _P]),
    "mpcb_last_transfer_bytes": (None, [C.c_void_p, C.POINTER(_I64), C.POINTER(_I64)]),
    "mpcb_plant_step": (C.c_int, [C.c_void_p, _I64, _P, _P, _P, _P, _P]),
    "mpcb_fp64_peak": (C.c_int, [C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "mpcb_latency_probe": (C.c_int, [C.POINTER(C.c_double)]),
    "mpcb_launch_count": (_I64, [C.c_void_p]),
    "mpcb_last_solve_bulk_profile": (None, [C.c_void_p, C.POINTER(C.c_double)]),
    "mpcb_last_solve_profile": (None, [C.c_void_p, C.POINTER(C.c_double)]),
}

_lib = None


class MpcbError(RuntimeError):
    pass


def lib():
    """Load libmpcb200.so once.  Raises if it has not been built (no fallback path exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MpcbError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a). mpcb200 has no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc, handle=None, what=""):
    if rc == 0:
        return
    msg = ""
    if handle:
        m = lib().mpcb_last_error(handle)
        msg = m.decode() if m else ""
    names = {E_ARG: "MPCB_E_ARG", E_CUDA: "MPCB_E_CUDA", E_UNSUPPORTED: "MPCB_E_UNSUPPORTED",
             E_NODEVICE: "MPCB_E_NODEVICE (no CUDA device; there is no CPU fallback)"}
    raise MpcbError(f"{what} failed: {names.get(rc, rc)} {msg}")


def default_desc(model_id=MODEL_PACEJKA6, N=12, S=100):
    d = MpcProblemDesc()
    lib().mpcb_default_desc(C.byref(d), model_id, N, S)
    return d


def default_solver_params(N=12):
    sp = MpcSolverParams()
    lib().mpcb_default_solver_params(C.byref(sp), N)
    return sp


def default_vehicle_params():
    import numpy as np
    p = (C.c_double * NP)()
    lib().mpcb_default_vehicle_params(p)
    return np.array(list(p), dtype=np.float64)
