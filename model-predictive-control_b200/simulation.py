"""Headless drop-in for reference simulation.py: same function names and arguments (simulation.py:5,48,
60,67,86).  matplotlib / a display are unavailable on this box, so every `plot_*` writes the arrays it
would have drawn to an .npz file (returned path) instead of opening a window.  `rotate_point` and the five-corner
car outline of `plot_car` are VERBATIM from reference simulation.py:60-83 (about 17 lines of plotting geometry: the
outline has to be the reference's for the saved arrays to be what its figures show).  Out of the hot path:
no GPU work here.
"""
import os
import re

import numpy as np

OUT_DIR = os.environ.get("MPCB200_PLOT_DIR", ".")


def _save(title, **arrays):
    name = re.sub(r"[^A-Za-z0-9_.-]+", "_", str(title)).strip("_") or "figure"
    path = os.path.join(OUT_DIR, name + ".npz")
    np.savez(path, **{k: np.asarray(v) for k, v in arrays.items()})
    return path


def plot_results(t, x, y, φ, vx, vy, ω, u, title):
    return _save(title, t=t, x=x, y=y, phi=φ, vx=vx, vy=vy, omega=ω, u=u)


def plot_trajectory(x, y, φ, u, title):
    δ = np.asarray(u)[1, :]
    return _save(title, x=x, y=y, heading=np.stack([np.cos(φ), np.sin(φ)]),
                 steer_dir=np.stack([np.cos(φ + δ), np.sin(φ + δ)]))


def rotate_point(point_x, point_y, origin_x, origin_y, angle):
    return [
        np.cos(angle) * (point_x - origin_x) - np.sin(angle) * (point_y - origin_y) + origin_x,
        np.sin(angle) * (point_x - origin_x) + np.cos(angle) * (point_y - origin_y) + origin_y
    ]


def car_outline(model, x, y, φ):
    """The five corner points reference plot_car draws (simulation.py:67-83)."""
    lf, lr = model.params.axis_front, model.params.axis_rear
    w = model.params.width
    x_left, x_right = x - lr, x + lf
    y_lower, y_upper = y - w / 2, y + w / 2
    return np.array([
        rotate_point(x_left, y_lower, x, y, φ),
        rotate_point(x_left, y_upper, x, y, φ),
        rotate_point(x_right, y_upper, x, y, φ),
        rotate_point(x_right, y_lower, x, y, φ),
        rotate_point(x_left, y_lower, x, y, φ)
    ])


def plot_car(model, x, y, φ):
    return car_outline(model, x, y, φ)


def simulate_motion(model, x, y, φ, vx, vy, u, t, title):
    """Animation frames as arrays: per tick the car outline, the front-axle point and steering direction."""
    x, y, φ = np.asarray(x), np.asarray(y), np.asarray(φ)
    δ = np.asarray(u)[1, :]
    outlines = np.stack([car_outline(model, x[i], y[i], φ[i]) for i in range(x.size)])
    front = np.array([rotate_point(x[i] + model.params.axis_front, y[i], x[i], y[i], φ[i]) for i in range(x.size)])
    return _save(title, outlines=outlines, front=front, steer_dir=np.stack([np.cos(φ + δ), np.sin(φ + δ)], 1),
                 x=x, y=y, vx=vx, t=t)
