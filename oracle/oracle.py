"""ctypes wrapper of the CPU oracle (oracle/libmpc_oracle.so) — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this.  "parity unpinned" against alpaqa/CasADi (not installable here); pinned against the reference's
NumPy code through tests/golden/reference_numpy.json (see tests/test_oracle_golden.py).
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
LIB_PATH = os.path.join(_HERE, "libmpc_oracle.so")

# struct layouts are shared with the product's header (include/mpcb200.h)
sys.path.insert(0, _ROOT)
import mpcb200  # noqa: E402
from mpcb200._ffi import MpcProblemDesc, MpcSolverParams  # noqa: E402

_lib = None
_D = C.c_double
_PD = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_PI = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")


FAST_LIB_PATH = os.path.join(_HERE, "libmpc_oracle_fast.so")


def build(force=False):
    if force or not os.path.exists(LIB_PATH) or not os.path.exists(FAST_LIB_PATH):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return LIB_PATH


def _bind(L):
    L.orc_wrap_to_pi.restype = _D
    L.orc_wrap_to_pi.argtypes = [_D]
    L.orc_u_trgl.restype = _D
    L.orc_u_trgl.argtypes = [_D, _D, _D]
    L.orc_u_rbd.restype = _D
    L.orc_u_rbd.argtypes = [_D, _D, _D, _D]
    L.orc_dnf.restype = _D
    L.orc_event_state.restype = _D
    L.orc_event_state.argtypes = [_D] * 5
    L.orc_num_threads.restype = C.c_int
    return L


class variant:
    """`with oracle.variant("fast"):` — route every call of this module through the second build of the same sources
    (oracle/Makefile: -O3, AVX2, fused multiply-adds).  Used only for the oracle-vs-oracle sensitivity yardstick."""

    def __init__(self, name):
        assert name == "fast"

    def __enter__(self):
        global _lib
        build()
        self._saved = lib()
        _lib = _bind(C.CDLL(FAST_LIB_PATH))

    def __exit__(self, *exc):
        global _lib
        _lib = self._saved


def lib():
    global _lib
    if _lib is None:
        build()
        L = _bind(C.CDLL(LIB_PATH))
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


DEFAULT_P = np.array([9.7e-2, 4.7e-2, 5e-2, 0.09, 0.07, 8e-2, 5.5e-2, 0.1735, 18.3e-5, 0.32, 1.0, 0.268, 2.165, 3.47,
                      0.242, 2.38, 2.84, 0.266, 0.1, 0.1025, 0.1629, 0.0011])  # main.py:82-111


def dynamics(model_id, x, u, p=DEFAULT_P, clip=False, acc=2.0, mu=1.0):
    x, u, p = f64(x), f64(u), f64(p)
    out = np.zeros(6)
    lib().orc_dynamics(C.c_int(model_id), _p(p), _D(acc), _D(mu), _p(x), _p(u), C.c_int(int(clip)), _p(out))
    return out[:4 if model_id == 1 else 6]


def dynamics_vjp(model_id, x, u, w, p=DEFAULT_P, acc=2.0, mu=1.0):
    xb, ub = np.zeros(6), np.zeros(2)
    x6, w6 = np.zeros(6), np.zeros(6)
    x6[:len(x)] = x
    w6[:len(w)] = w
    lib().orc_dynamics_vjp(C.c_int(model_id), _p(f64(p)), _D(acc), _D(mu), _p(x6), _p(f64(u)), _p(w6), _p(xb), _p(ub))
    return xb[:len(x)], ub


def rk4_stage(desc, x, u, p=DEFAULT_P):
    out = np.zeros(6)
    x6 = np.zeros(6)
    x6[:len(x)] = x
    lib().orc_rk4_stage(C.byref(desc), _p(f64(p)), _p(x6), _p(f64(u)), _p(out))
    return out[:len(x)]


def wrap_to_pi(a):
    return lib().orc_wrap_to_pi(float(a))


def road_errors(road_flat, pos, heading, normalised=False):
    road = f64(road_flat)
    err = np.zeros(3)
    idx = lib().orc_road_errors(C.c_int(road.size // 2), _p(road), _D(pos[0]), _D(pos[1]), _D(heading),
                                C.c_int(int(normalised)), _p(err))
    return int(idx), err


def bezier_sample(P12, S):
    out = np.zeros(2 * S)
    lib().orc_bezier_sample(_p(f64(P12).ravel()), C.c_int(S), _p(out))
    return out


def bezier_point(P12, j):
    xy = np.zeros(2)
    lib().orc_bezier_point(_p(f64(P12).ravel()), _D(j), _p(xy))
    return xy


def bezier_control_points(i, h, W, theta, l, Lf, v0, v1, D1):
    P, tca = np.zeros(12), _D()
    lib().orc_bezier_control_points(_D(i), _D(h), _D(W), _D(theta), _D(l), _D(Lf), _D(v0), _D(v1), _D(D1), _p(P),
                                    C.byref(tca))
    return P.reshape(2, 6), tca.value


def safe_distances(X_e, X_obs, desc=None):
    """dpf_test.py:57-64 get_safe_distances(X_e, X_obs) -> (x_s, y_s)."""
    d = default_desc() if desc is None else desc
    out = np.zeros(2)
    lib().orc_safe_distances(C.byref(d), _p(f64(X_e)), _p(f64(X_obs)), _p(out))
    return out[0], out[1]


def rows_per_stage(desc):
    nx = 4 if desc.model_id == 1 else 6
    return nx + (desc.n_obs if desc.flags & 16 else 0)


def u_trgl(y, y_t, a=0.5):
    return lib().orc_u_trgl(y, y_t, a)


def u_rbd(y, y_br, y_bl, b=100.0):
    return lib().orc_u_rbd(y, y_br, y_bl, b)


def dnf(x, y, phi, v, n_obs, xo, yo, po, vo, Af=1000.0, b=1.0, sx=2.0, sy=0.5):
    xo, yo, po, vo = f64(xo), f64(yo), f64(po), f64(vo)
    return lib().orc_dnf(_D(x), _D(y), _D(phi), _D(v), C.c_int(n_obs), _p(xo), _p(yo), _p(po), _p(vo), _D(Af), _D(b),
                         _D(sx), _D(sy))


def event_state(x, ux, a, t0, t1):
    return lib().orc_event_state(x, ux, a, t0, t1)


def _i32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.int32)


def rollout(desc, x0, U, road, p=DEFAULT_P, road_idx=None, obs=None):
    nx = 4 if desc.model_id == 1 else 6
    x0 = f64(x0).reshape(-1, nx)
    B, N = x0.shape[0], desc.N
    U, road, p = f64(U).reshape(B, 2 * N), f64(road), f64(p)
    obs = None if obs is None else f64(obs)
    road_idx = _i32(road_idx)
    states, idx, cost = np.zeros((B, N, nx)), np.zeros((B, N), np.int32), np.zeros(B)
    lib().orc_rollout_batch(C.byref(desc), C.c_int64(B), _p(x0), _p(U), _p(p), _p(road), _p(road_idx), _p(obs),
                            _p(states), _p(idx), _p(cost))
    return states, idx, cost


def value_grad(desc, x0, U, road, p=DEFAULT_P, road_idx=None, obs=None, lam=None, Sigma=None):
    nx = 4 if desc.model_id == 1 else 6
    x0 = f64(x0).reshape(-1, nx)
    B, N = x0.shape[0], desc.N
    U, road, p = f64(U).reshape(B, 2 * N), f64(road), f64(p)
    obs = None if obs is None else f64(obs)
    lam = None if lam is None else f64(lam)
    Sigma = None if Sigma is None else f64(Sigma)
    road_idx = _i32(road_idx)
    cost, grad = np.zeros(B), np.zeros((B, 2 * N))
    lib().orc_value_grad_batch(C.byref(desc), C.c_int64(B), _p(x0), _p(U), _p(p), _p(road), _p(road_idx), _p(obs),
                               _p(lam), _p(Sigma), _p(cost), _p(grad))
    return cost, grad


def solve(desc, sp, x0, road, U, lam=None, p=DEFAULT_P, road_idx=None, obs=None, lb=None, ub=None, threads=0,
          want_times=False):
    """Returns dict(U, lam, status, outer, inner, evals_value, evals_grad, psi, res[, times])."""
    nx = 4 if desc.model_id == 1 else 6
    x0 = f64(x0).reshape(-1, nx)
    B, N = x0.shape[0], desc.N
    U = f64(U).reshape(B, 2 * N).copy()
    lam = None if lam is None else f64(lam).reshape(B, rows_per_stage(desc) * N).copy()
    road, p = f64(road), f64(p)
    obs = None if obs is None else f64(obs)
    road_idx = _i32(road_idx)
    if lb is None:
        lb = np.tile([-p[10], -p[9]], N)
        ub = -lb
    lb, ub = f64(lb), f64(ub)
    status, iters, evals = np.zeros(B, np.int32), np.zeros((B, 2), np.int32), np.zeros((B, 2), np.int32)
    psi, res = np.zeros(B), np.zeros(B)
    times = np.zeros(B) if want_times else None
    lib().orc_solve_batch(C.byref(desc), C.byref(sp), C.c_int64(B), _p(x0), _p(p), _p(road), _p(road_idx), _p(obs),
                          _p(lb), _p(ub), _p(U), _p(lam), _p(status), _p(iters), _p(evals), _p(psi), _p(res),
                          _p(times), C.c_int(threads))
    out = dict(U=U, lam=lam, status=status, outer=iters[:, 0], inner=iters[:, 1], evals_value=evals[:, 0],
               evals_grad=evals[:, 1], psi=psi, res=res)
    if want_times:
        out["times"] = times
    return out


def fb_residual(desc, x0, U, road, gamma, p=DEFAULT_P, road_idx=None, obs=None, lb=None, ub=None):
    """(r [B], psi [B]): r = ||Pi_C(U - gamma grad psi(U)) - U||_2 with the oracle's gradient, psi = psi(U)."""
    nx = 4 if desc.model_id == 1 else 6
    x0 = f64(x0).reshape(-1, nx)
    B, N = x0.shape[0], desc.N
    U, road, p, gamma = f64(U).reshape(B, 2 * N), f64(road), f64(p), f64(gamma).reshape(B)
    obs = None if obs is None else f64(obs)
    if lb is None:
        lb = np.tile([-p[10], -p[9]], N)
        ub = -lb
    lb, ub = f64(lb), f64(ub)
    res, psi = np.zeros(B), np.zeros(B)
    lib().orc_fb_residual_batch(C.byref(desc), C.c_int64(B), _p(x0), _p(U), _p(p), _p(road), _p(_i32(road_idx)), _p(obs),
                                _p(lb), _p(ub), _p(gamma), _p(res), _p(psi))
    return res, psi


def game_payoff(ego, cars, car_par=None):
    """game_theory.py:205-244 batched: ego [B,3], cars [B,M,3] as (x, v, lane; lane 0 = no car in that slot) ->
    dict(payoff [B,2], decision [B], ctrl [B,12], tca [B])."""
    ego = f64(ego).reshape(-1, 3)
    B = ego.shape[0]
    cars = f64(cars).reshape(B, -1, 3)
    M = cars.shape[1]
    payoff, decision = np.zeros((B, 2)), np.zeros(B, np.int32)
    ctrl, tca = np.zeros((B, 12)), np.zeros(B)
    lib().orc_game_payoff_batch(C.c_int64(B), C.c_int(M), _p(ego), _p(cars), _p(None if car_par is None else f64(car_par)),
                                _p(payoff), _p(decision), _p(ctrl), _p(tca))
    return dict(payoff=payoff, decision=decision, ctrl=ctrl, tca=tca)


def count_ops(desc, x0, U, road, p=DEFAULT_P, obs=None):
    counts = (C.c_longlong * 4)()
    lib().orc_count_ops(C.byref(desc), _p(f64(x0)), _p(f64(U)), _p(f64(p)), _p(f64(road)),
                        _p(None if obs is None else f64(obs)), counts)
    return dict(grad_arith=counts[0], grad_transc=counts[1], value_arith=counts[2], value_transc=counts[3])


def num_threads():
    return lib().orc_num_threads()


# ---- host-side copies of the defaults (the oracle must not need the CUDA library) ------------------
def default_desc(model_id=0, N=12, S=100):
    d = MpcProblemDesc()
    d.model_id, d.N, d.S, d.n_sub = model_id, N, S, 4
    d.road_mode, d.n_roads, d.flags, d.n_obs = 0, 1, 0, 0
    d.Ts, d.v_ref = 0.05, 1.0
    for i, v in enumerate([0.5, 1, 1, 0.5, 0.1, 0.01]):
        d.c[i] = v
    d.kin_friction, d.kin_accel = 1.0, 2.0
    d.lane_a, d.lane_y_t, d.lane_b, d.lane_y_br, d.lane_y_bl = 0.5, 1.75, 100.0, 1.0, 6.0
    d.obs_Af, d.obs_b, d.obs_sigma_x, d.obs_sigma_y, d.obs_alpha_div = 1000.0, 1.0, 2.0, 0.5, 5.0
    for i, v in enumerate([20, 1, 1, 2, 1, 0.1]):
        d.g_offset[i] = v
        d.D_lower[i] = -np.inf
        d.D_upper[i] = np.inf
    d.safe_x0, d.safe_y0, d.safe_ax_max, d.safe_ay_max = 5.0, 3.0, 3.0, 1.0   # dpf_test.py:6
    return d


def default_solver_params(N=12):
    sp = MpcSolverParams()
    sp.max_iter, sp.lbfgs_mem, sp.max_no_progress, sp.hess_vec_fd = 1000, N, 10, 1
    sp.Lgamma_factor, sp.L_min, sp.L_max, sp.lip_eps, sp.lip_delta = 0.95, 1e-5, 1e20, 1e-6, 1e-12
    eps = np.finfo(np.float64).eps
    sp.tau_min, sp.quad_tol, sp.ls_tol = 1.0 / 256, 10 * eps, 10 * eps
    sp.alm_max_iter, sp.eps_target, sp.delta_target, sp.Sigma0 = 1000, 1e-6, 1e-4, 1e5
    sp.eps0, sp.rho, sp.theta, sp.Delta, sp.M, sp.Sigma_max = 1.0, 0.1, 0.1, 10.0, 1e9, 1e9
    return sp
