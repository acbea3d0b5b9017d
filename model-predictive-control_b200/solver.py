"""Batched MPC problem on the GPU: torch tensors for memory/streams, libmpcb200.so for every kernel.

`BatchedMpc` is the object behind the reference-facing shims (controller.py, car_dynamics.py, main.py):
it owns one `MpcHandle` (the "compiled problem" that CasADi codegen produces in reference main.py:54)
and exposes the hot-path operations on `[B, ...]` fp64 CUDA tensors.
"""
import ctypes as C
from dataclasses import dataclass

import numpy as np
import torch

from . import _ffi
from ._ffi import (FLAG_BOUNDED_D, FLAG_EXACT, FLAG_LANE, FLAG_OBST, FLAG_SAFE_DIST, MODEL_KINEMATIC4, MODEL_PACEJKA6, ROAD_BEZIER,
                   ROAD_PER_INSTANCE, ROAD_SHARED, ROAD_TABLE)


@dataclass
class SolveStats:
    status: torch.Tensor       # [B] int32, alpaqa SolverStatus order
    outer_iterations: torch.Tensor  # [B] int32
    inner_iterations: torch.Tensor  # [B] int32
    evals_value: torch.Tensor  # [B] int32  psi-only evaluations
    evals_grad: torch.Tensor   # [B] int32  psi + grad psi evaluations
    psi: torch.Tensor          # [B] f64
    residual: torch.Tensor     # [B] f64 final ||p||_2


def _dev_ptr(t, dtype, what):
    if t is None:
        return None
    if not (isinstance(t, torch.Tensor) and t.is_cuda):
        raise TypeError(f"{what} must be a CUDA tensor (mpcb200 has no CPU path)")
    if t.dtype != dtype or not t.is_contiguous():
        raise TypeError(f"{what} must be contiguous {dtype}, got {t.dtype} contiguous={t.is_contiguous()}")
    return C.c_void_p(t.data_ptr())


def _on_device(fn):
    """Run a BatchedMpc method with the handle's device current (launches and allocations of the library land on the
    current device; the caller's current device may be another GPU)."""
    import functools

    @functools.wraps(fn)
    def wrapper(self, *args, **kw):
        with torch.cuda.device(self.device):
            return fn(self, *args, **kw)
    return wrapper


class BatchedMpc:
    """One problem description (model, horizon, road layout, cost terms) bound to the current CUDA device."""

    def __init__(self, N=12, S=100, model_id=MODEL_PACEJKA6, road_mode=ROAD_SHARED, n_roads=1, flags=0, n_obs=0,
                 Ts=0.05, v_ref=1.0, c=None, n_sub=4, D_lower=None, D_upper=None, desc_overrides=None,
                 device=None):
        if not torch.cuda.is_available():
            raise _ffi.MpcbError("no CUDA device: mpcb200 has no CPU fallback")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self._lib = _ffi.lib()
        d = _ffi.default_desc(model_id, N, S)
        d.road_mode, d.n_roads, d.flags, d.n_obs = road_mode, n_roads, flags, n_obs
        d.Ts, d.v_ref, d.n_sub = Ts, v_ref, n_sub
        if c is not None:
            for i in range(6):
                d.c[i] = float(c[i])
        if D_lower is not None:
            for i, v in enumerate(D_lower):
                d.D_lower[i] = float(v)
        if D_upper is not None:
            for i, v in enumerate(D_upper):
                d.D_upper[i] = float(v)
        for k, v in (desc_overrides or {}).items():
            setattr(d, k, v)
        self.desc = d
        self.N, self.S, self.model_id = N, S, model_id
        self.nx = 4 if model_id == MODEL_KINEMATIC4 else 6
        # rows of g per stage: the nx state rows (main.py:46-51) + one safe-distance ellipse per obstacle (MPCB_FLAG_SAFE_DIST)
        self.rows = self.nx + (n_obs if flags & FLAG_SAFE_DIST else 0)
        self.n, self.m = 2 * N, self.rows * N
        self._h = C.c_void_p()
        with torch.cuda.device(self.device):
            _ffi.check(self._lib.mpcb_create(C.byref(d), C.byref(self._h)), None, "mpcb_create")
        self.p = torch.tensor(_ffi.default_vehicle_params(), device=self.device)  # main.py:82-111

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self._lib.mpcb_destroy(h)
            except Exception:
                pass

    # ---- helpers ------------------------------------------------------------------------------
    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _t(self, x, shape=None, dtype=torch.float64):
        # numpy / python inputs are uploaded (the reference-facing convenience); a torch tensor is taken as is:
        # the wrong device or dtype is a loud error, never a silent conversion or a CPU path
        if isinstance(x, torch.Tensor) and (x.device != self.device or x.dtype != dtype):
            raise TypeError(f"expected a {dtype} tensor on {self.device}, got {x.dtype} on {x.device}")
        t = torch.as_tensor(x, dtype=dtype, device=self.device).contiguous()
        if shape is not None:
            t = t.reshape(shape)
        return t

    def launch_count(self):
        return int(self._lib.mpcb_launch_count(self._h))

    def last_solve_profile(self):
        """dict(mode, tile, rounds, eval_ms, trans_ms, eval_launches) of the last solve (CUDA events inside the library)."""
        out = (C.c_double * 6)()
        self._lib.mpcb_last_solve_profile(self._h, out)
        prof = dict(mode="rounds" if out[0] else "persistent", tile=int(out[1]), rounds=int(out[2]), eval_ms=out[3],
                    trans_ms=out[4], eval_launches=int(out[5]))
        if out[0]:
            b = (C.c_double * 5)()
            self._lib.mpcb_last_solve_bulk_profile(self._h, b)
            prof["bulk"] = dict(rounds=int(b[0]), eval_ms=b[1], trans_ms=b[2], grad_requests=b[3], value_requests=b[4])
        return prof

    def default_solver_params(self):
        return _ffi.default_solver_params(self.N)

    def default_box(self):
        """C = [-max_drive, max_drive] x [-max_steer, max_steer] tiled (main.py:55-56)."""
        p = self.p.cpu().numpy()
        lb = np.tile([-p[10], -p[9]], self.N)
        return self._t(lb), self._t(-lb)

    # ---- hot-path operations --------------------------------------------------------------------
    @_on_device
    def dynamics(self, x, u, clip=False, p=None):
        x, u = self._t(x).reshape(-1, self.nx), self._t(u).reshape(-1, 2)
        out = torch.empty_like(x)
        p = self.p if p is None else self._t(p)
        _ffi.check(self._lib.mpcb_dynamics(self._h, x.shape[0], _dev_ptr(x, torch.float64, "x"),
                                           _dev_ptr(u, torch.float64, "u"), _dev_ptr(p, torch.float64, "p"),
                                           int(clip), _dev_ptr(out, torch.float64, "out"), self._stream()),
                   self._h, "mpcb_dynamics")
        return out

    @_on_device
    def plant_step(self, x, u, p=None):
        x, u = self._t(x).reshape(-1, self.nx), self._t(u).reshape(-1, 2)
        out = torch.empty_like(x)
        p = self.p if p is None else self._t(p)
        _ffi.check(self._lib.mpcb_plant_step(self._h, x.shape[0], _dev_ptr(x, torch.float64, "x"),
                                             _dev_ptr(u, torch.float64, "u"), _dev_ptr(p, torch.float64, "p"),
                                             _dev_ptr(out, torch.float64, "out"), self._stream()),
                   self._h, "mpcb_plant_step")
        return out

    @_on_device
    def road_errors(self, pos, heading, road, road_idx=None, normalised=False):
        pos, heading, road = self._t(pos).reshape(-1, 2), self._t(heading).reshape(-1), self._t(road)
        B = pos.shape[0]
        road_idx = None if road_idx is None else self._t(road_idx, dtype=torch.int32)
        idx = torch.empty(B, dtype=torch.int32, device=self.device)
        err = torch.empty(B, 3, dtype=torch.float64, device=self.device)
        _ffi.check(self._lib.mpcb_road_errors(self._h, B, _dev_ptr(pos, torch.float64, "pos"),
                                              _dev_ptr(heading, torch.float64, "heading"),
                                              _dev_ptr(road, torch.float64, "road"),
                                              _dev_ptr(road_idx, torch.int32, "road_idx"), int(normalised),
                                              _dev_ptr(idx, torch.int32, "idx"), _dev_ptr(err, torch.float64, "err"),
                                              self._stream()), self._h, "mpcb_road_errors")
        return idx, err

    @_on_device
    def bezier_sample(self, ctrl):
        ctrl = self._t(ctrl).reshape(-1, 12)
        out = torch.empty(ctrl.shape[0], 2 * self.S, dtype=torch.float64, device=self.device)
        _ffi.check(self._lib.mpcb_bezier_sample(self._h, ctrl.shape[0], _dev_ptr(ctrl, torch.float64, "ctrl"),
                                                _dev_ptr(out, torch.float64, "out"), self._stream()),
                   self._h, "mpcb_bezier_sample")
        return out

    def _common(self, x0, road, road_idx, obs, p):
        x0 = self._t(x0).reshape(-1, self.nx)
        road = self._t(road)
        road_idx = None if road_idx is None else self._t(road_idx, dtype=torch.int32)
        obs = None if obs is None else self._t(obs)
        p = self.p if p is None else self._t(p)
        return x0, road, road_idx, obs, p

    @_on_device
    def rollout(self, x0, U, road, road_idx=None, obs=None, p=None):
        """model.simulate + cost sum: returns (states [B,N,nx], nearest idx [B,N], cost [B])."""
        x0, road, road_idx, obs, p = self._common(x0, road, road_idx, obs, p)
        B = x0.shape[0]
        U = self._t(U).reshape(B, self.n)
        states = torch.empty(B, self.N, self.nx, dtype=torch.float64, device=self.device)
        idx = torch.empty(B, self.N, dtype=torch.int32, device=self.device)
        cost = torch.empty(B, dtype=torch.float64, device=self.device)
        f64 = torch.float64
        _ffi.check(self._lib.mpcb_rollout(self._h, B, _dev_ptr(x0, f64, "x0"), _dev_ptr(U, f64, "U"),
                                          _dev_ptr(p, f64, "p"), _dev_ptr(road, f64, "road"),
                                          _dev_ptr(road_idx, torch.int32, "road_idx"), _dev_ptr(obs, f64, "obs"),
                                          _dev_ptr(states, f64, "states"), _dev_ptr(idx, torch.int32, "idx"),
                                          _dev_ptr(cost, f64, "cost"), self._stream()), self._h, "mpcb_rollout")
        return states, idx, cost

    @_on_device
    def value_grad(self, x0, U, road, road_idx=None, obs=None, lam=None, Sigma=None, p=None, out=None):
        """psi and grad_U psi (generated-C f + grad_f)."""
        x0, road, road_idx, obs, p = self._common(x0, road, road_idx, obs, p)
        B = x0.shape[0]
        U = self._t(U).reshape(B, self.n)
        lam = None if lam is None else self._t(lam).reshape(B, self.m)
        Sigma = None if Sigma is None else self._t(Sigma).reshape(B, self.m)
        if out is None:
            cost = torch.empty(B, dtype=torch.float64, device=self.device)
            grad = torch.empty(B, self.n, dtype=torch.float64, device=self.device)
        else:
            cost, grad = out
        f64 = torch.float64
        _ffi.check(self._lib.mpcb_value_grad(self._h, B, _dev_ptr(x0, f64, "x0"), _dev_ptr(U, f64, "U"),
                                             _dev_ptr(p, f64, "p"), _dev_ptr(road, f64, "road"),
                                             _dev_ptr(road_idx, torch.int32, "road_idx"), _dev_ptr(obs, f64, "obs"),
                                             _dev_ptr(lam, f64, "lam"), _dev_ptr(Sigma, f64, "Sigma"),
                                             _dev_ptr(cost, f64, "cost"), _dev_ptr(grad, f64, "grad"),
                                             self._stream()), self._h, "mpcb_value_grad")
        return cost, grad

    def value_grad_static(self, x0, U, road, cost, grad, road_idx=None, obs=None, lam=None, Sigma=None, p=None):
        """mpcb_value_grad_static: psi and grad psi with NO allocation anywhere (static device scratch; see
        include/mpcb200.h for the limits).  Every argument must already be a contiguous CUDA tensor of the right
        dtype on this device — nothing is converted, copied or allocated here — and `cost` [B], `grad` [B, 2N]
        are written in place."""
        f64, i32 = torch.float64, torch.int32
        p = self.p if p is None else p
        B = x0.shape[0]
        with torch.cuda.device(self.device):
            _ffi.check(self._lib.mpcb_value_grad_static(
                self._h, B, _dev_ptr(x0, f64, "x0"), _dev_ptr(U, f64, "U"), _dev_ptr(p, f64, "p"),
                _dev_ptr(road, f64, "road"), _dev_ptr(road_idx, i32, "road_idx"), _dev_ptr(obs, f64, "obs"),
                _dev_ptr(lam, f64, "lam"), _dev_ptr(Sigma, f64, "Sigma"), _dev_ptr(cost, f64, "cost"),
                _dev_ptr(grad, f64, "grad"), self._stream()), self._h, "mpcb_value_grad_static")
        return cost, grad

    @_on_device
    def solve(self, x0, road, U, lam=None, road_idx=None, obs=None, lb=None, ub=None, sp=None, p=None):
        """The MPC step's NLP solve (controller.py:57).  U (and lam) are updated IN PLACE (warm start
        in, solution out) and also returned together with per-instance statistics."""
        x0, road, road_idx, obs, p = self._common(x0, road, road_idx, obs, p)
        B = x0.shape[0]
        if not (isinstance(U, torch.Tensor) and U.is_cuda and U.dtype == torch.float64 and U.is_contiguous()):
            U = self._t(U).reshape(B, self.n).clone()
        if lam is not None and not (isinstance(lam, torch.Tensor) and lam.is_cuda and lam.is_contiguous()):
            lam = self._t(lam).reshape(B, self.m).clone()
        if lb is None or ub is None:
            lb, ub = self.default_box()
        lb, ub = self._t(lb), self._t(ub)
        sp = self.default_solver_params() if sp is None else sp
        dev, i32, f64 = self.device, torch.int32, torch.float64
        status = torch.empty(B, dtype=i32, device=dev)
        iters = torch.empty(B, 2, dtype=i32, device=dev)
        evals = torch.empty(B, 2, dtype=i32, device=dev)
        psi = torch.empty(B, dtype=f64, device=dev)
        res = torch.empty(B, dtype=f64, device=dev)
        _ffi.check(self._lib.mpcb_solve(self._h, C.byref(sp), B, _dev_ptr(x0, f64, "x0"), _dev_ptr(p, f64, "p"),
                                        _dev_ptr(road, f64, "road"), _dev_ptr(road_idx, i32, "road_idx"),
                                        _dev_ptr(obs, f64, "obs"), _dev_ptr(lb, f64, "lb"), _dev_ptr(ub, f64, "ub"),
                                        _dev_ptr(U, f64, "U"), _dev_ptr(lam, f64, "lam"),
                                        _dev_ptr(status, i32, "status"), _dev_ptr(iters, i32, "iters"),
                                        _dev_ptr(evals, i32, "evals"), _dev_ptr(psi, f64, "psi"),
                                        _dev_ptr(res, f64, "res"), self._stream()), self._h, "mpcb_solve")
        return U, lam, SolveStats(status, iters[:, 0], iters[:, 1], evals[:, 0], evals[:, 1], psi, res)

    def solve_host(self, x0, road, U, lam=None, road_idx=None, obs=None, lb=None, ub=None, sp=None, p=None):
        """Same solve with numpy (host) buffers: pinned staging + H2D + solve + D2H inside the library.
        U / lam are updated in place when they are contiguous float64 arrays."""
        f8 = np.float64
        x0 = np.ascontiguousarray(x0, f8).reshape(-1, self.nx)
        B = x0.shape[0]
        road = np.ascontiguousarray(road, f8)
        U = np.ascontiguousarray(U, f8).reshape(B, self.n)
        lam = None if lam is None else np.ascontiguousarray(lam, f8).reshape(B, self.m)
        road_idx = None if road_idx is None else np.ascontiguousarray(road_idx, np.int32)
        obs = None if obs is None else np.ascontiguousarray(obs, f8)
        p = _ffi.default_vehicle_params() if p is None else np.ascontiguousarray(p, f8)
        if lb is None or ub is None:
            lb = np.tile([-p[10], -p[9]], self.N).astype(f8)
            ub = -lb
        lb, ub = np.ascontiguousarray(lb, f8), np.ascontiguousarray(ub, f8)
        sp = self.default_solver_params() if sp is None else sp
        status = np.empty(B, np.int32)
        iters = np.empty((B, 2), np.int32)
        evals = np.empty((B, 2), np.int32)
        psi, res = np.empty(B, f8), np.empty(B, f8)

        def hp(a):
            return None if a is None else C.c_void_p(a.ctypes.data)

        with torch.cuda.device(self.device):
            _ffi.check(self._lib.mpcb_solve_host(self._h, C.byref(sp), B, hp(x0), hp(p), hp(road), hp(road_idx),
                                                 hp(obs), hp(lb), hp(ub), hp(U), hp(lam), hp(status), hp(iters),
                                                 hp(evals), hp(psi), hp(res)), self._h, "mpcb_solve_host")
        return U, lam, SolveStats(status, iters[:, 0], iters[:, 1], evals[:, 0], evals[:, 1], psi, res)

    @_on_device
    def game_payoff(self, ego, cars, car_params=None):
        """game_theory.py:205-244 for a batch: ego [B,3], cars [B,M,3] of (x, v, lane) -> (payoff [B,2], decision [B] int32,
        ctrl [B,12], tca [B]) on the device (mpcb_game_payoff)."""
        ego = self._t(ego).reshape(-1, 3)
        B = ego.shape[0]
        cars = self._t(cars).reshape(B, -1, 3)
        M = cars.shape[1]
        f64 = torch.float64
        payoff = torch.empty(B, 2, dtype=f64, device=self.device)
        decision = torch.empty(B, dtype=torch.int32, device=self.device)
        ctrl = torch.empty(B, 12, dtype=f64, device=self.device)
        tca = torch.empty(B, dtype=f64, device=self.device)
        cp = None if car_params is None else np.ascontiguousarray(car_params, np.float64)
        _ffi.check(self._lib.mpcb_game_payoff(self._h, B, M, _dev_ptr(ego, f64, "ego"), _dev_ptr(cars, f64, "cars"),
                                              None if cp is None else C.c_void_p(cp.ctypes.data),
                                              _dev_ptr(payoff, f64, "payoff"), _dev_ptr(decision, torch.int32, "decision"),
                                              _dev_ptr(ctrl, f64, "ctrl"), _dev_ptr(tca, f64, "tca"), self._stream()),
                   self._h, "mpcb_game_payoff")
        return payoff, decision, ctrl, tca

    def check_guards(self):
        """Canary bytes around the handle's scratch buffers that were overwritten (handle created under MPCB_GUARDS=1)."""
        return int(self._lib.mpcb_check_guards(self._h))

    def last_solve_latency_ms(self, B):
        """[B] milliseconds from the start of the last solve on the GPU to the moment each instance returned its answer
        (GPU %globaltimer stamps written by the kernels)."""
        out = torch.empty(B + 1, dtype=torch.int64, device=self.device)
        with torch.cuda.device(self.device):
            _ffi.check(self._lib.mpcb_last_finish_ns(self._h, B, C.c_void_p(out.data_ptr()), self._stream()),
                       self._h, "mpcb_last_finish_ns")
        return (out[:B] - out[B]).double() * 1e-6

    def set_trace(self, cap):
        """Record psi at the top of every inner iteration of later solves (0 = off): the reference's progress callback."""
        _ffi.check(self._lib.mpcb_set_trace(self._h, int(cap)), self._h, "mpcb_set_trace")
        self._trace_cap = int(cap)

    def last_trace(self, B):
        """[B, cap, 4] (psi, phi_gamma, gamma, accepted candidate) per inner iteration of the last solve (NaN = not reached)."""
        cap = getattr(self, "_trace_cap", 0)
        out = torch.empty(B, cap, 4, dtype=torch.float64, device=self.device)
        with torch.cuda.device(self.device):
            _ffi.check(self._lib.mpcb_last_trace(self._h, B, _dev_ptr(out, torch.float64, "trace"), self._stream()),
                       self._h, "mpcb_last_trace")
        return out

    def last_inner_failures(self, B):
        """[B] int32: PANOC runs of the last solve's ALM loops that did not converge (alpaqa: inner_convergence_failures)."""
        out = torch.empty(B, dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            _ffi.check(self._lib.mpcb_last_inner_failures(self._h, B, _dev_ptr(out, torch.int32, "out"), self._stream()),
                       self._h, "mpcb_last_inner_failures")
        return out

    def last_final_gamma(self, B):
        """[B] final PANOC step sizes of the last solve on this handle (alpaqa's inner stats `final γ`)."""
        out = torch.empty(B, dtype=torch.float64, device=self.device)
        with torch.cuda.device(self.device):
            _ffi.check(self._lib.mpcb_last_final_gamma(self._h, B, _dev_ptr(out, torch.float64, "gamma"), self._stream()),
                       self._h, "mpcb_last_final_gamma")
        return out

    def last_transfer_bytes(self):
        a, b = C.c_int64(), C.c_int64()
        self._lib.mpcb_last_transfer_bytes(self._h, C.byref(a), C.byref(b))
        return a.value, b.value


def fp64_peak_tflops():
    """Measured dependent-DFMA-chain throughput of the current device (roofline denominator)."""
    t, ms = C.c_double(), C.c_double()
    _ffi.check(_ffi.lib().mpcb_fp64_peak(C.byref(t), C.byref(ms)), None, "mpcb_fp64_peak")
    return t.value, ms.value
