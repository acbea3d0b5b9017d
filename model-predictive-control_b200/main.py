"""Drop-in for reference main.py: `get_centerline`, `create_casadi_problem`, `alpaqa_vehicle_test`.

`create_casadi_problem` keeps its signature (main.py:25) but returns an mpcb200 problem instead of
CasADi-generated C: there is nothing to code-generate, the kernels are compiled once into libmpcb200.so.
"""
import numpy as np

from . import alpaqa_compat as pa
from .car_dynamics import KinematicBicyclePacejka
from .controller import MPCController
from .workloads import get_centerline  # main.py:11-22

__all__ = ["get_centerline", "create_casadi_problem", "alpaqa_vehicle_test", "VEHICLE_PARAM"]

# main.py:82-111
VEHICLE_PARAM = np.array([9.7e-2, 4.7e-2, 5e-2, 0.09, 0.07, 8e-2, 5.5e-2, 0.1735, 18.3e-5, 0.32, 1.0,
                          0.268, 2.165, 3.47, 0.242, 2.38, 2.84, 0.266, 0.1, 0.1025, 0.1629, 0.0011])


def create_casadi_problem(model, N_horiz, centerline_size, v_ref, max_drive, max_steer):
    """main.py:25-59: cost = sum of stage costs over the RK4 rollout, g = 6N state rows, box C on the
    inputs, D left unbounded (main.py:57 is commented out in the reference)."""
    prob = pa.Problem(N_horiz, centerline_size, model_id=getattr(model, "model_id", 0), v_ref=v_ref)
    prob.C.lowerbound = np.tile([-max_drive, -max_steer], N_horiz)
    prob.C.upperbound = np.tile([max_drive, max_steer], N_horiz)
    return prob


def alpaqa_vehicle_test(N_sim=400, N_horiz=12, centerline_size=100, v_ref=1.0, is_straight=True, verbose=False):
    """main.py:62-167 without the plotting: closed loop of N_sim ticks (solve -> apply first input ->
    plant RK4 step).  Returns (y_mpc [6, N_sim], controller)."""
    model = KinematicBicyclePacejka()
    model.dynamics()
    y_n = np.array([0, 0, 0, .5, 0, 0], dtype=np.float64)      # main.py:72-79
    max_drive, max_steer = 1.0, 0.32
    param = VEHICLE_PARAM
    centerline_val = get_centerline(centerline_size, is_straight=is_straight).ravel(order='F')
    prob = create_casadi_problem(model, N_horiz, centerline_size, v_ref, max_drive, max_steer)
    prob.param = np.concatenate((y_n, centerline_val, param))   # main.py:119
    controller = MPCController(model, prob, N_horiz, verbose=verbose)
    y_mpc = np.empty((y_n.shape[0], N_sim))
    for n in range(N_sim):
        U = controller(y_n, centerline_val)                     # main.py:140
        u_n = model.input_to_matrix(U)[:, 0]
        y_n = model.simulate(1, y_n, u_n, param).T.ravel()      # main.py:145
        y_mpc[:, n] = y_n
    if verbose:
        print(controller.tot_it, controller.failures)
    return y_mpc, controller


if __name__ == '__main__':
    y, ctl = alpaqa_vehicle_test(verbose=True)
    np.savez("alpaqa_vehicle_test.npz", y_mpc=y)
