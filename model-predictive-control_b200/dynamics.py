"""Drop-in for reference dynamics.py: `VehicleParameters`, `KinematicBicyclePacejka`,
`KinematicBicycleSimplified` with the same `__call__(x, u, t=None) -> xdot` (dynamics.py:67,144),
evaluated by the CUDA library (mpcb_dynamics, clip=1 mirrors clip_inputs at dynamics.py:99/166).
`x` may be (nx,) or (B, nx); `u` (2,) or (B, 2).
"""
import numpy as np

from . import _ffi


class VehicleParameters:  # dynamics.py:4-42 (same names, same values)
    length: float = 9.7e-2
    axis_front: float = 4.7e-2
    axis_rear: float = 5e-2
    front: float = 0.09
    rear: float = 0.07
    width: float = 8e-2
    height: float = 5.5e-2
    mass: float = 0.1735
    inertia: float = 18.3e-5
    max_steer: float = 0.32
    max_drive: float = 1.0
    bf: float = 0.268
    cf: float = 2.165
    df: float = 3.47
    br: float = 0.242
    cr: float = 2.38
    dr: float = 2.84
    friction: float = 1
    acceleration: float = 2
    cm1: float = 0.266
    cm2: float = 0.1
    cr0: float = 0.1025
    cr1: float = 0.1629
    cr2: float = 0.0011

    def as_vector(self):
        """The 22-vector in car_dynamics.py:65-88 order."""
        return np.array([self.length, self.axis_front, self.axis_rear, self.front, self.rear, self.width, self.height,
                         self.mass, self.inertia, self.max_steer, self.max_drive, self.bf, self.cf, self.df, self.br,
                         self.cr, self.dr, self.cm1, self.cm2, self.cr0, self.cr1, self.cr2], dtype=np.float64)


class _Model:
    model_id = _ffi.MODEL_PACEJKA6

    def __init__(self, params: VehicleParameters = None) -> None:
        self.params = VehicleParameters() if params is None else params
        self._prob = None

    def clip_inputs(self, d=None, δ=None):  # dynamics.py:56-65
        res = []
        if d is not None:
            res.append(np.clip(d, -self.params.max_drive, self.params.max_drive))
        if δ is not None:
            res.append(np.clip(δ, -self.params.max_steer, self.params.max_steer))
        if len(res) == 1:
            return res[0]
        return tuple(res)

    def __call__(self, x, u, t=None):
        from .solver import BatchedMpc
        if self._prob is None:
            self._prob = BatchedMpc(N=1, S=3, model_id=self.model_id, desc_overrides={
                "kin_friction": float(self.params.friction), "kin_accel": float(self.params.acceleration)})
        x, u = np.asarray(x, dtype=np.float64), np.asarray(u, dtype=np.float64)
        out = self._prob.dynamics(np.atleast_2d(x), np.atleast_2d(u), clip=True, p=self.params.as_vector()).cpu().numpy()
        return out[0] if x.ndim == 1 else out


class KinematicBicyclePacejka(_Model):
    model_id = _ffi.MODEL_PACEJKA6


class KinematicBicycleSimplified(_Model):
    model_id = _ffi.MODEL_KINEMATIC4
