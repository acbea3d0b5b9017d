"""Drop-in for reference car_dynamics.py: same class and method names, numeric instead of CasADi-symbolic.

The reference builds SX graphs that CasADi compiles to C (car_dynamics.py:93-147,230-258); here every
method evaluates on the GPU through libmpcb200.so.  `X`, `u`, `params` keep their *shapes* (the only
property callers use: main.py:27,29).
"""
import numpy as np

from . import _ffi


class _Sym:
    """Stand-in for a CasADi SX column: only `.shape` is consumed by callers."""

    def __init__(self, n):
        self.shape = (n, 1)


class KinematicBicyclePacejka:
    model_id = _ffi.MODEL_PACEJKA6

    def __init__(self):
        self.X, self.u, self.params = _Sym(6), _Sym(2), _Sym(_ffi.NP)
        self.f = None
        self.f_d = None
        self.Ts = 0.05
        self._mpc = {}

    def _prob(self, N=1, S=100, Ts=None):
        from .solver import BatchedMpc
        key = (N, S, self.Ts if Ts is None else Ts)
        if key not in self._mpc:
            self._mpc[key] = BatchedMpc(N=N, S=S, model_id=self.model_id, road_mode=_ffi.ROAD_PER_INSTANCE, Ts=key[2])
        return self._mpc[key]

    def dynamics(self, Ts=0.05):
        """car_dynamics.py:93-147: returns f_d(y, u, p) -> y+ (RK4, 4 finite elements over Ts); also sets f."""
        self.Ts = Ts
        prob = self._prob(1, 100, Ts)
        self.f = lambda y, u, p=None: prob.dynamics(np.atleast_2d(y), np.atleast_2d(u), clip=False, p=p).cpu().numpy()
        self.f_d = lambda y, u, p=None: prob.plant_step(np.atleast_2d(y), np.atleast_2d(u), p=p).cpu().numpy()
        return self.f_d

    def input_to_matrix(self, u):
        """car_dynamics.py:149-156: flat [d0, δ0, d1, δ1, ...] -> 2 x N (column-major)."""
        u = np.asarray(u)
        return u.reshape((2, u.shape[0] // 2), order='F')

    def simulate(self, N_sim, y_0, u, p=None):
        """car_dynamics.py:158-166 (mapaccum): 6 x N_sim matrix, column n = state AFTER input n."""
        u = np.asarray(u, dtype=np.float64)
        U = u.reshape(2, -1, order='F') if u.ndim == 1 else u
        prob = self._prob(N_sim, 3)
        dummy_road = np.zeros((1, 6))
        st, _, _ = prob.rollout(np.asarray(y_0, float).reshape(1, -1), U.T.reshape(1, -1), dummy_road, p=p)
        return st.cpu().numpy()[0].T

    def wrap_to_pi(self, angle):
        return np.mod(angle + np.pi, 2 * np.pi) - np.pi  # car_dynamics.py:168-172 verbatim maths

    def _errors(self, size, pos, heading, flat):
        prob = self._prob(1, size)
        return prob.road_errors(np.asarray(pos, float).reshape(1, 2), np.array([float(heading)]),
                                np.asarray(flat, float).reshape(1, -1), normalised=False)

    def find_nearest_point(self, size, vehicle_position, centerline):
        """car_dynamics.py:174-192: (nearest, previous, next) vertices; centerline is (size, 2)."""
        c = np.asarray(centerline, float).reshape(size, 2)
        idx, _ = self._errors(size, vehicle_position, 0.0, c.ravel(order='F'))
        i = int(idx[0])
        return c[i], c[max(i - 1, 0)], c[i + 1]

    def compute_errors(self, size, vehicle_position, vehicle_heading, centerline_flat):
        """car_dynamics.py:194-228: (cte, heading_error, pos_error), raw cross products."""
        _, err = self._errors(size, vehicle_position, vehicle_heading, centerline_flat)
        e = err.cpu().numpy()[0]
        return e[0], e[1], e[2]

    def generate_stage_cost_fun(self, centerline_size, target_v, c=np.array([0.5, 1, 1, .5, 0.1, 0.01])):
        """car_dynamics.py:230-258: returns L_cost(X, u, centerline)."""
        def L_cost(X, u, centerline):
            X, u = np.asarray(X, float).ravel(), np.asarray(u, float).ravel()
            cte, he, pe = self.compute_errors(centerline_size, X[:2], X[2], centerline)
            return (c[0] * (np.sqrt(X[3] ** 2 + X[4] ** 2) - target_v) ** 2 + c[1] * cte ** 2 + c[2] * pe ** 2 +
                    c[3] * he ** 2 + c[4] * u[1] ** 2 + c[5] * u[0] ** 2)
        return L_cost
