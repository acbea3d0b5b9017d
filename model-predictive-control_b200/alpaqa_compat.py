"""The slice of the alpaqa 0.0.1 Python API that reference controller.py / main.py touch, backed by
libmpcb200.so.  `import mpcb200.alpaqa_compat as pa` makes controller.py's code read as
in the reference:

    inner = pa.StructuredPANOCLBFGSSolver(panoc_params={...}, lbfgs_params={'memory': N})
    solver = pa.ALMSolver(alm_params={...}, inner_solver=inner)
    x, y, stats = solver(problem, x, y)          # controller.py:57

Differences, all deliberate: wall-clock caps ('max_time', controller.py:30,44) are accepted and ignored
(they make results time-dependent); the per-iteration Python progress callback (controller.py:23-24,38)
is accepted and never called (no host round trip inside the fused kernel); x / y / problem.param may
carry a leading batch dimension.
"""
import enum
import time
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


class StructuredPANOCLBFGSProgressInfo:  # type named by controller.py:23,49; never instantiated here
    pass


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

    def mpc(self):
        """The device-side problem (created on first use; rebuilt if D's finiteness pattern changes)."""
        from .solver import BatchedMpc
        Dl = np.asarray(self.D.lowerbound).reshape(self.N_horiz, self.nx)[0]
        Du = np.asarray(self.D.upperbound).reshape(self.N_horiz, self.nx)[0]
        bounded = bool(np.isfinite(Dl).any() or np.isfinite(Du).any())
        key = (bounded, tuple(Dl), tuple(Du))
        if self._mpc is None or key != self._mpc_key:
            self._mpc = BatchedMpc(N=self.N_horiz, S=self.centerline_size, model_id=self.model_id,
                                   road_mode=_ffi.ROAD_PER_INSTANCE, v_ref=self.v_ref, c=self.c, Ts=self.Ts,
                                   flags=self.flags | (_ffi.FLAG_BOUNDED_D if bounded else 0), n_obs=self.n_obs,
                                   D_lower=Dl, D_upper=Du)
            self._mpc_key = key
        return self._mpc


class StructuredPANOCLBFGSSolver:
    def __init__(self, panoc_params=None, lbfgs_params=None):
        self.panoc_params = dict(panoc_params or {})
        self.lbfgs_params = dict(lbfgs_params or {})
        self._callback = None

    def set_progress_callback(self, cb):
        self._callback = cb  # kept for API parity; the fused kernel never calls back to the host

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
        # 'max_time' (controller.py:30), 'hessian_step_size_heuristic' (controller.py:32; rule not recoverable,
        # SURVEY Appendix A) and 'print_interval' are accepted and ignored.
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
        if par.shape[0] == 1 and B > 1:
            par = np.broadcast_to(par, (B, par.shape[1]))
        y0 = np.ascontiguousarray(par[:, :nx])
        road = np.ascontiguousarray(par[:, nx:nx + 2 * S])
        p = np.ascontiguousarray(par[0, nx + 2 * S:nx + 2 * S + _ffi.NP])
        lam = None
        if y is not None:
            lam = np.ascontiguousarray(np.atleast_2d(np.asarray(y, dtype=np.float64))).copy()
        mpc = problem.mpc()
        sp = self._params(problem.N_horiz)
        X, lam, st = mpc.solve_host(y0, road, X, lam=lam, lb=np.asarray(problem.C.lowerbound, float),
                                    ub=np.asarray(problem.C.upperbound, float), sp=sp, p=p)
        elapsed = timedelta(seconds=time.perf_counter() - t0)
        status = [SolverStatus(int(s)) for s in st.status]
        if batched:
            stats = {"status": status, "outer_iterations": st.outer_iterations.copy(),
                     "inner": {"iterations": st.inner_iterations.copy()}, "elapsed_time": elapsed,
                     "inner_convergence_failures": np.zeros(B, int), "ε": st.residual.copy(), "ψ": st.psi.copy()}
            return X, lam, stats
        stats = {"status": status[0], "outer_iterations": int(st.outer_iterations[0]),
                 "inner": {"iterations": int(st.inner_iterations[0])}, "elapsed_time": elapsed,
                 "inner_convergence_failures": 0, "ε": float(st.residual[0]), "ψ": float(st.psi[0])}
        return X[0], (None if lam is None else lam[0]), stats
