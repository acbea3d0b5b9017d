"""The slice of the alpaqa 0.0.1 Python API that reference controller.py / main.py touch, backed by
libmpcb200.so.  `import mpcb200.alpaqa_compat as pa` makes controller.py's code read as
in the reference:

    inner = pa.StructuredPANOCLBFGSSolver(panoc_params={...}, lbfgs_params={'memory': N})
    solver = pa.ALMSolver(alm_params={...}, inner_solver=inner)
    x, y, stats = solver(problem, x, y)          # controller.py:57

Differences, all deliberate: wall-clock caps ('max_time', controller.py:30,44) are accepted and ignored
(they make results time-dependent); 'hessian_step_size_heuristic' (controller.py:32) is accepted, ignored and warned
about once (alpaqa 0.0.1's rule is not recoverable here: SURVEY Appendix A 3c); the per-iteration Python progress
callback (controller.py:23-24,38) is fed AFTER the solve from a per-iteration trace the kernel records (no host round
trip inside the solver), for instance 0 of a batch; x / y / problem.param may carry a leading batch dimension.
"""
import enum
import time
import warnings
from datetime import timedelta

import numpy as np

from . import _ffi


class SolverStatus(enum.IntEnum):  # alpaqa 0.0.1 order
    Unknown = 0
    Converged = 1
    MaxTime = 2
    MaxIter = 3
    NotFinite = 4
    NoProgress = 5
    Interrupted = 6


class PANOCStopCrit(enum.Enum):
    ApproxKKT = 0
    ProjGradNorm = 1
    ProjGradNorm2 = 2
    ProjGradUnitNorm = 3
    FPRNorm = 4


ProjGradNorm2 = PANOCStopCrit.ProjGradNorm2


class StructuredPANOCLBFGSProgressInfo:
    """What the reference's callback receives (controller.py:23-24 reads `.ψ`): one per inner iteration, built from
    the kernel's trace buffer after the solve."""

    def __init__(self, k=0, ψ=float("nan"), φγ=float("nan"), γ=float("nan"), instance=0):
        self.k, self.ψ, self.φγ, self.γ, self.instance = k, ψ, φγ, γ, instance


class Box:
    def __init__(self, n):
        self.lowerbound = np.full(n, -np.inf)
        self.upperbound = np.full(n, np.inf)


class Problem:
    """What `generate_and_compile_casadi_problem` returns in the reference (main.py:54): n, m, writable
    `param`, boxes C and D — plus the description the CUDA library needs instead of generated C."""

    def __init__(self, N_horiz, centerline_size, model_id=_ffi.MODEL_PACEJKA6, v_ref=1.0, c=None, Ts=0.05,
                 flags=0, n_obs=0):
        self.N_horiz, self.centerline_size, self.model_id = N_horiz, centerline_size, model_id
        self.nx = 4 if model_id == _ffi.MODEL_KINEMATIC4 else 6
        self.n, self.m = 2 * N_horiz, self.nx * N_horiz
        self.v_ref, self.c, self.Ts, self.flags, self.n_obs = v_ref, c, Ts, flags, n_obs
        self.param = np.zeros(self.nx + 2 * centerline_size + _ffi.NP)  # [y0, centerline, p] main.py:30,119
        self.C, self.D = Box(self.n), Box(self.m)
        self._mpc = None
        self._mpc_key = None

    def mpc(self, road_mode=_ffi.ROAD_SHARED, n_roads=1):
        """The device-side problem for one road layout (created on first use; rebuilt if D's finiteness pattern
        changes).  The reference's only case — ONE centerline in `param` — is MPCB_ROAD_SHARED: the road is staged in
        shared memory and the nearest-vertex scan is pruned.  A batch over a few distinct centerlines becomes a road
        table; only a batch of many distinct centerlines pays for the per-instance layout."""
        from .solver import BatchedMpc
        Dl = np.asarray(self.D.lowerbound).reshape(self.N_horiz, self.nx)[0]
        Du = np.asarray(self.D.upperbound).reshape(self.N_horiz, self.nx)[0]
        bounded = bool(np.isfinite(Dl).any() or np.isfinite(Du).any())
        key = (bounded, tuple(Dl), tuple(Du), road_mode, n_roads)
        if self._mpc is None:
            self._mpc = {}
        if key not in self._mpc:
            self._mpc[key] = BatchedMpc(N=self.N_horiz, S=self.centerline_size, model_id=self.model_id,
                                        road_mode=road_mode, n_roads=n_roads, v_ref=self.v_ref, c=self.c, Ts=self.Ts,
                                        flags=self.flags | (_ffi.FLAG_BOUNDED_D if bounded else 0), n_obs=self.n_obs,
                                        D_lower=Dl, D_upper=Du)
        return self._mpc[key]


MAX_TABLE_ROADS = 32


def dedupe_roads(road, max_roads=MAX_TABLE_ROADS):
    """(table [R, 2S], idx [B] int32) if the rows of `road` [B, 2S] take at most `max_roads` distinct values, else
    None.  One vectorised pass over the whole array per distinct road, no gathered copies (the reference's callers pass
    one road, or a handful): 45 ms for 65 536 x 200 doubles drawn from two roads (three times that with `road[left]`)."""
    B = road.shape[0]
    idx = np.full(B, -1, np.int32)
    table = []
    while True:
        left = np.flatnonzero(idx < 0)
        if left.size == 0:
            break
        if len(table) == max_roads:
            return None
        r = road[left[0]]
        same = np.equal(road, r).all(axis=1)
        same &= idx < 0
        idx[same] = len(table)
        table.append(r)
    return np.ascontiguousarray(np.stack(table)), idx


_warned_heuristic = False


class StructuredPANOCLBFGSSolver:
    def __init__(self, panoc_params=None, lbfgs_params=None):
        self.panoc_params = dict(panoc_params or {})
        self.lbfgs_params = dict(lbfgs_params or {})
        self._callback = None

    def set_progress_callback(self, cb):
        self._callback = cb  # replayed after the solve from the kernel's per-iteration trace (ALMSolver.__call__)

    def fill(self, sp):
        pp = self.panoc_params
        if "max_iter" in pp:
            sp.max_iter = int(pp["max_iter"])
        if "stop_crit" in pp and pp["stop_crit"] not in (ProjGradNorm2, "ProjGradNorm2"):
            raise NotImplementedError("only stop_crit=ProjGradNorm2 (reference controller.py:29) is implemented")
        if "memory" in self.lbfgs_params:
            sp.lbfgs_mem = int(self.lbfgs_params["memory"])
        for k_py, k_c in (("τ_min", "tau_min"), ("L_min", "L_min"), ("L_max", "L_max"), ("Lγ_factor", "Lgamma_factor"),
                          ("max_no_progress", "max_no_progress")):
            if k_py in pp:
                setattr(sp, k_c, type(getattr(sp, k_c))(pp[k_py]))
        # 'max_time' (controller.py:30) and 'print_interval' are accepted and ignored.
        global _warned_heuristic
        if pp.get("hessian_step_size_heuristic", 0) and not _warned_heuristic:
            _warned_heuristic = True
            warnings.warn("mpcb200: panoc_params['hessian_step_size_heuristic'] is ignored (alpaqa 0.0.1's step-size rule is "
                          "not recoverable; the canonical solver of SURVEY Appendix A runs without it): iteration counts "
                          "can differ from alpaqa's", stacklevel=3)
        return sp


class ALMSolver:
    def __init__(self, alm_params=None, inner_solver=None):
        self.alm_params = dict(alm_params or {})
        self.inner_solver = inner_solver or StructuredPANOCLBFGSSolver()

    def _params(self, N):
        sp = _ffi.default_solver_params(N)
        self.inner_solver.fill(sp)
        ap = self.alm_params
        for k_py, k_c in (("ε", "eps_target"), ("δ", "delta_target"), ("Σ_0", "Sigma0"), ("ε_0", "eps0"), ("ρ", "rho"),
                          ("θ", "theta"), ("Δ", "Delta"), ("M", "M"), ("Σ_max", "Sigma_max")):
            if k_py in ap:
                setattr(sp, k_c, float(ap[k_py]))
        if "max_iter" in ap:
            sp.alm_max_iter = int(ap["max_iter"])
        return sp

    def __call__(self, problem, x, y):
        """solver(problem, x, y) -> (x, y, stats)  (controller.py:57).  Batched when x is (B, n)."""
        t0 = time.perf_counter()
        x = np.asarray(x, dtype=np.float64)
        batched = x.ndim == 2
        X = np.ascontiguousarray(np.atleast_2d(x)).copy()
        B, nx, S = X.shape[0], problem.nx, problem.centerline_size
        par = np.atleast_2d(np.asarray(problem.param, dtype=np.float64))
        y0 = par[:, :nx]
        if y0.shape[0] == 1 and B > 1:
            y0 = np.broadcast_to(y0, (B, nx))
        y0 = np.ascontiguousarray(y0)
        p = np.ascontiguousarray(par[0, nx + 2 * S:nx + 2 * S + _ffi.NP])
        # road layout: one centerline for everybody (the reference's case) / a few / one per instance
        roads = par[:, nx:nx + 2 * S]
        road_idx = None
        if roads.shape[0] == 1:
            mpc, road = problem.mpc(_ffi.ROAD_SHARED), np.ascontiguousarray(roads[0])
        else:
            dd = dedupe_roads(roads)
            if dd is None:
                mpc, road = problem.mpc(_ffi.ROAD_PER_INSTANCE), np.ascontiguousarray(roads)
            elif dd[0].shape[0] == 1:
                mpc, road = problem.mpc(_ffi.ROAD_SHARED), dd[0][0]
            else:
                mpc, road, road_idx = problem.mpc(_ffi.ROAD_TABLE, dd[0].shape[0]), dd[0], dd[1]
        lam = None
        if y is not None:
            lam = np.ascontiguousarray(np.atleast_2d(np.asarray(y, dtype=np.float64))).copy()
        sp = self._params(problem.N_horiz)
        cb = self.inner_solver._callback
        cap = 0
        if cb is not None:  # per-iteration psi of instance 0, replayed to the callback after the solve
            cap = int(min(max(sp.max_iter, 1) * 8, 1 << 14))
        mpc.set_trace(cap)
        X, lam, st = mpc.solve_host(y0, road, X, lam=lam, road_idx=road_idx, lb=np.asarray(problem.C.lowerbound, float),
                                    ub=np.asarray(problem.C.upperbound, float), sp=sp, p=p)
        ifail = mpc.last_inner_failures(B).cpu().numpy()
        if cb is not None:
            tr = mpc.last_trace(1)[0].cpu().numpy()
            for k, v in enumerate(tr[:int(st.inner_iterations[0]) + int(st.outer_iterations[0])]):
                if v[0] == v[0]:
                    cb(StructuredPANOCLBFGSProgressInfo(k=k, ψ=float(v[0]), φγ=float(v[1]), γ=float(v[2]), instance=0))
        elapsed = timedelta(seconds=time.perf_counter() - t0)
        status = [SolverStatus(int(s)) for s in st.status]
        extra = {"hessian_step_size_heuristic": "off (ignored)", "road_layout": {0: "shared", 1: "table", 2: "per_instance"}[mpc.desc.road_mode]}
        if batched:
            stats = {"status": status, "outer_iterations": st.outer_iterations.copy(),
                     "inner": {"iterations": st.inner_iterations.copy()}, "elapsed_time": elapsed,
                     "inner_convergence_failures": ifail.astype(int), "ε": st.residual.copy(), "ψ": st.psi.copy(), **extra}
            return X, lam, stats
        stats = {"status": status[0], "outer_iterations": int(st.outer_iterations[0]),
                 "inner": {"iterations": int(st.inner_iterations[0])}, "elapsed_time": elapsed,
                 "inner_convergence_failures": int(ifail[0]), "ε": float(st.residual[0]), "ψ": float(st.psi[0]), **extra}
        return X[0], (None if lam is None else lam[0]), stats
