"""mpcb200 — B200-native batched nonlinear MPC (ALM + structured PANOC + L-BFGS).

Drop-in for the one hot path of panagiotou23/model-predictive-control: the NLP solve inside
MPCController.__call__ (reference controller.py:51-69).  Import as `mpcb200` (the root-level shim
`mpcb200.py` loads this directory, whose on-disk name contains a hyphen, under that name).

Importing the package never touches CUDA; the shared library is loaded on first use and every
compute entry point raises if it is missing or no GPU is present (no CPU fallback).
"""
__version__ = "0.1.0"

from . import _ffi  # noqa: F401
from ._ffi import (FLAG_BOUNDED_D, FLAG_EXACT, FLAG_LANE, FLAG_OBST, FLAG_SAFE_DIST, MODEL_KINEMATIC4, MODEL_PACEJKA6, ROAD_BEZIER,  # noqa: F401
                   ROAD_PER_INSTANCE, ROAD_SHARED, ROAD_TABLE, MpcbError)


def __getattr__(name):  # lazy: solver needs torch
    if name in ("BatchedMpc", "SolveStats", "fp64_peak_tflops"):
        from . import solver
        return getattr(solver, name)
    raise AttributeError(name)
