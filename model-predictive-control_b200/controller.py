"""Drop-in for reference controller.py: same class, constructor, call signature and public attributes
(controller.py:10-69); the alpaqa objects are mpcb200.alpaqa_compat's, which run the fused CUDA solver.

Batched extension (backwards compatible): `y_n` may be (B, 6) and `centerline` (2S,) shared
or (B, 2S); `__call__` then returns (B, 2N) and `U` / `λ` carry a leading batch dimension.
"""
from datetime import timedelta

import numpy as np

from . import alpaqa_compat as pa


class MPCController:

    def __init__(self, model, problem, N_horiz, verbose=False):
        self.model = model
        self.problem = problem

        self.N_horiz = 12 if N_horiz is None else N_horiz
        self.u_dim = 2
        self.tot_it = 0
        self.failures = 0
        self.verbose = verbose
        self.U = np.tile([1.0, 0.0], self.N_horiz)                    # controller.py:20
        self.λ = np.zeros((problem.nx * self.N_horiz,))              # controller.py:21

        def progress_callback(info):  # controller.py:23-24 (never invoked: no host callback in the kernel)
            print("cost: ", info.ψ)

        inner_solver = pa.StructuredPANOCLBFGSSolver(
            panoc_params={
                'stop_crit': pa.ProjGradNorm2,
                'max_time': timedelta(seconds=0.2),   # accepted, ignored (results must not depend on time)
                'max_iter': 1000,
                'hessian_step_size_heuristic': 15,
            },
            lbfgs_params={'memory': self.N_horiz},
        )
        inner_solver.set_progress_callback(progress_callback)
        self.solver = pa.ALMSolver(
            alm_params={
                'ε': 1e-6,
                'δ': 1e-4,
                'Σ_0': 1e5,
                'max_time': timedelta(seconds=0.5),   # accepted, ignored
                'max_iter': 1000,
            },
            inner_solver=inner_solver,
        )
        self.info = pa.StructuredPANOCLBFGSProgressInfo
        self.last_stats = None

    def __call__(self, y_n, centerline):
        y_n = np.asarray(y_n, dtype=np.float64)
        centerline = np.asarray(centerline, dtype=np.float64)
        nx = self.problem.nx
        if y_n.ndim == 2 and y_n.shape[1] == nx and y_n.shape[0] != 1 or (y_n.ndim == 2 and centerline.ndim == 2):
            return self._call_batched(y_n, centerline)
        y_n = y_n.ravel()
        # Set the current state as the initial state (controller.py:54)
        self.problem.param[:(y_n.shape[0] + centerline.shape[0])] = np.concatenate((y_n, centerline))
        # Solve, warm-starting from the previous solution and multipliers — no horizon shift (controller.py:57)
        self.U, self.λ, stats = self.solver(self.problem, self.U, self.λ)
        if self.verbose:
            print(stats['status'], stats['outer_iterations'], stats['inner']['iterations'], stats['elapsed_time'],
                  stats['inner_convergence_failures'])
        self.tot_it += stats['inner']['iterations']
        self.failures += stats['status'] != pa.SolverStatus.Converged
        self.last_stats = stats
        return self.U

    def _call_batched(self, y_n, centerline):
        B = y_n.shape[0]
        S2 = centerline.shape[-1]
        par = np.empty((B, self.problem.param.shape[-1]))
        par[:] = np.atleast_2d(self.problem.param)[-1]
        par[:, :y_n.shape[1]] = y_n
        par[:, y_n.shape[1]:y_n.shape[1] + S2] = centerline
        self.problem.param = par
        if self.U.ndim == 1:
            self.U = np.tile(self.U, (B, 1))
            self.λ = np.tile(self.λ, (B, 1))
        self.U, self.λ, stats = self.solver(self.problem, self.U, self.λ)
        self.tot_it += int(np.sum(stats['inner']['iterations']))
        self.failures += int(sum(s != pa.SolverStatus.Converged for s in stats['status']))
        self.last_stats = stats
        return self.U


def mpc_controller(model, current_state, road, target_velocity, N=2, dt=0.1):
    """reference controller.py:72-108 (legacy): explicit-Euler rollout of `model(x, u, t)` over N stages and
    `scipy.optimize.minimize` (BFGS, finite-difference gradient) on cte / heading / speed penalties; returns the first
    input.  Kept importable with the reference's semantics; NOT on the accelerated path (superseded by the
    alpaqa controller).  `model` and `road` are the reference-named objects of mpcb200.dynamics / mpcb200.road, whose
    evaluations run on the GPU.  One deliberate difference: the reference passes `np.array(x[0], x[1])` as the
    position (controller.py:89), which raises a TypeError in NumPy; the position is built as `np.array([x[0], x[1]])`."""
    from scipy.optimize import minimize

    def cost_fn(u_flat, *args):
        states, N, dt, current_state, road, target_velocity = args
        cost = 0
        u = u_flat.reshape(2, N)
        cte_weight, heading_weight, velocity_weight = 100, 10, 10
        for i in range(N):
            if i == 0:
                x = np.copy(current_state)
            else:
                x = x + model(states[i - 1], u[:, i - 1], dt * (i - 1)) * dt
            states[i] = x
            _, heading_error, cte = road.compute_errors(np.array([x[0], x[1]]), x[2])
            velocity = np.sqrt(x[3] ** 2 + x[4] ** 2)
            cost += cte_weight * cte ** 2 + heading_weight * heading_error ** 2 + velocity_weight * velocity
        return cost

    states = [None] * N
    u_flat = np.zeros((N * 2,))
    result = minimize(cost_fn, u_flat, args=(states, N, dt, np.asarray(current_state, dtype=np.float64), road, target_velocity))
    return result.x.reshape(2, N)[:, 0]


def mpc_controller_initial(model, current_state, target_state, N=2, dt=0.1):
    """reference controller.py:111-135 (legacy): same rollout, cost = sum of squared distances to `target_state`."""
    from scipy.optimize import minimize
    target_state = np.asarray(target_state, dtype=np.float64)

    def cost_fn(u_flat, *args):
        states, target_state, N, dt = args
        cost = 0
        u = u_flat.reshape(2, N)
        for i in range(N):
            if i == 0:
                x = np.copy(current_state)
            else:
                x = x + model(states[i - 1], u[:, i - 1], dt * (i - 1)) * dt
            states[i] = x
            cost += np.sum((x - target_state) ** 2)
        return cost

    states = [None] * N
    u_flat = np.zeros((N * 2,))
    result = minimize(cost_fn, u_flat, args=(states, target_state, N, dt))
    return result.x.reshape(2, N)[:, 0]


def straight_line_controller(current_state, road):
    """reference controller.py:138-156 (debug stub): constant full drive, no steering."""
    return np.array([1, 0])
