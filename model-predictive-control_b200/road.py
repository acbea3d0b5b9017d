"""Drop-in for reference road.py: `Road(center=None)` with `wrap_to_pi`, `find_nearest_point`,
`compute_errors` (road.py:4-81) — NumPy semantics (plain argmin, errors divided by the segment length),
evaluated by the CUDA library (mpcb_road_errors, normalised=1).  Positions may be (2,) or (B, 2).
"""
import numpy as np

from . import _ffi


class Road:
    def __init__(self, center=None) -> None:
        if center is None:  # road.py:10-18: circle of radius 5 centred at (0, 5), 100 points
            theta = np.linspace(0, 2 * np.pi, 100)
            x = 5 * np.cos(theta)
            y = 5 * np.sin(theta) + 5
            self.centerline = np.stack((x, y), axis=1)
        else:
            self.centerline = np.asarray(center, dtype=np.float64)
        self._prob = None

    def wrap_to_pi(self, angle):
        return np.mod(angle + np.pi, 2 * np.pi) - np.pi

    def _query(self, pos, heading):
        from .solver import BatchedMpc
        if self._prob is None or self._prob.S != len(self.centerline):
            self._prob = BatchedMpc(N=1, S=len(self.centerline))
        pos = np.atleast_2d(np.asarray(pos, dtype=np.float64))
        heading = np.broadcast_to(np.asarray(heading, dtype=np.float64), (pos.shape[0],)).copy()
        idx, err = self._prob.road_errors(pos, heading, self.centerline.ravel(order="F"), normalised=True)
        return idx.cpu().numpy(), err.cpu().numpy()

    def find_nearest_point(self, vehicle_position=None):
        idx, _ = self._query(vehicle_position, 0.0)
        if np.ndim(vehicle_position) == 1:
            return int(idx[0]), self.centerline[int(idx[0])]
        return idx, self.centerline[idx]

    def compute_errors(self, vehicle_position, vehicle_heading):
        """-> (cte, heading_error, pos_error)   road.py:52-81"""
        _, err = self._query(vehicle_position, vehicle_heading)
        if np.ndim(vehicle_position) == 1:
            return err[0, 0], err[0, 1], err[0, 2]
        return err[:, 0], err[:, 1], err[:, 2]
