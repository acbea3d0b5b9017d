"""Generate tests/golden/reference_numpy.json by IMPORTING the reference's own NumPy code.

Run in the build container only (needs /root/reference, which does not exist on the GPU box):

    python tests/golden/gen_golden.py

The reference's CasADi/alpaqa path cannot run here, so these vectors pin the pieces of
the hot path whose reference code does run: dynamics.py (both ODEs), road.py (nearest vertex, errors,
wrap_to_pi), bezier_curves.py, dpf_test.py potentials, event_triggered.py's displacement formula and
game_theory.py's Bezier control points.  matplotlib is absent, so a stub module is installed first.
"""
import json
import os
import sys
import types

import numpy as np

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_numpy.json")


def _stub_matplotlib():
    class _Any(types.ModuleType):
        def __getattr__(self, name):
            if name.startswith("__"):
                raise AttributeError(name)
            return _Any(name)

        def __call__(self, *a, **k):
            return _Any("obj")

        def __iter__(self):
            return iter([_Any("a"), _Any("b")])

    for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.cm"):
        sys.modules[name] = _Any(name)


def main():
    _stub_matplotlib()
    sys.path.insert(0, REF)
    import dynamics  # noqa
    import road as road_mod  # noqa
    import bezier_curves  # noqa
    import dpf_test  # noqa
    import game_theory  # noqa

    rng = np.random.Generator(np.random.PCG64(20261104))
    out = {"generator": "tests/golden/gen_golden.py", "numpy": np.__version__}

    # ---- dynamics.py:67 Pacejka-6 (inputs are clipped inside, dynamics.py:99)
    pac = dynamics.KinematicBicyclePacejka()
    cases = [([0, 0, 0, .5, 0, 0], [.3, .1]), ([1, -.2, .3, 1.2, .05, -.4], [-.5, -.2]),
             ([0, 0, 0, .5, 0, 0], [2.0, 0.5])]
    for _ in range(40):
        x = [rng.uniform(-5, 5), rng.uniform(-5, 5), rng.uniform(-4, 4), rng.uniform(0.3, 2.0),
             rng.uniform(-0.3, 0.3), rng.uniform(-2, 2)]
        u = [rng.uniform(-1.3, 1.3), rng.uniform(-0.45, 0.45)]
        cases.append((x, u))
    for _ in range(6):  # reversing / negative vx exercises sign(vx) and atan2 quadrants
        x = [0.0, 0.0, rng.uniform(-3, 3), rng.uniform(-1.0, -0.3), rng.uniform(-0.2, 0.2), rng.uniform(-1, 1)]
        cases.append((x, [rng.uniform(-1, 1), rng.uniform(-0.3, 0.3)]))
    out["pacejka"] = [{"x": list(map(float, x)), "u": list(map(float, u)),
                       "xdot": pac(np.array(x, float), np.array(u, float)).tolist()} for x, u in cases]

    # ---- dynamics.py:144 Kinematic-4
    kin = dynamics.KinematicBicycleSimplified()
    cases = [([0, 0, .1, 1], [.3, .1]), ([2, 1, -.7, .4], [-1, .32])]
    for _ in range(30):
        cases.append(([rng.uniform(-5, 5), rng.uniform(-5, 5), rng.uniform(-4, 4), rng.uniform(0, 3)],
                      [rng.uniform(-1.3, 1.3), rng.uniform(-0.45, 0.45)]))
    out["kinematic"] = [{"x": list(map(float, x)), "u": list(map(float, u)),
                         "xdot": kin(np.array(x, float), np.array(u, float)).tolist()} for x, u in cases]

    # ---- road.py: default circle and the straight line of main.py:13
    def errs(rd, pos, heading):
        idx, _ = rd.find_nearest_point(np.array(pos, float))
        cte, he, pe = rd.compute_errors(np.array(pos, float), heading)
        return {"pos": list(map(float, pos)), "heading": float(heading), "idx": int(idx),
                "cte": float(cte), "heading_err": float(he), "pos_err": float(pe)}

    circle = road_mod.Road()
    straight_pts = np.array([[i / 10 - 0.1, 0] for i in range(100)])
    straight = road_mod.Road(straight_pts)
    rc = [errs(circle, p, h) for p, h in [((4.0, 7.5), 2.0), ((-3.0, 1.2), -1.0), ((0.2, 9.7), 3.0)]]
    for _ in range(40):
        th = rng.uniform(0.2, 6.0)
        r = 5 + rng.uniform(-0.4, 0.4)
        pos = (r * np.cos(th), r * np.sin(th) + 5)
        e = errs(circle, pos, rng.uniform(-3.5, 3.5))
        if 0 < e["idx"] < 98:
            rc.append(e)
    rs = [errs(straight, p, h) for p, h in [((0.33, 0.05), 0.1), ((5.02, -0.2), -0.3)]]
    for _ in range(40):
        pos = (rng.uniform(0.05, 9.6), rng.uniform(-0.4, 0.4))
        e = errs(straight, pos, rng.uniform(-3.5, 3.5))
        if 0 < e["idx"] < 98:
            rs.append(e)
    out["road_circle"] = {"centerline_flat_F": circle.centerline.ravel(order="F").tolist(), "cases": rc}
    out["road_straight"] = {"centerline_flat_F": straight_pts.ravel(order="F").tolist(), "cases": rs}
    out["wrap_to_pi"] = [{"a": float(a), "w": float(circle.wrap_to_pi(a))}
                         for a in [3.5, -3.5, 7.0, np.pi, -np.pi, 0.0, 1e-3, -12.0, 25.0] +
                         list(rng.uniform(-20, 20, 20))]

    # ---- bezier_curves.py
    bz = []
    for i in range(1, 11):
        P, tca = bezier_curves.get_bezier_control_points(i)
        pts = [list(map(float, bezier_curves.bezier_curve(j, P))) for j in np.linspace(0, 1, 11)]
        bz.append({"i": i, "P": P.tolist(), "tca": float(tca), "j": np.linspace(0, 1, 11).tolist(), "pts": pts})
    out["bezier"] = {"consts": {"h": bezier_curves.h, "W": bezier_curves.W, "theta": float(bezier_curves.θ),
                                "l": bezier_curves.l, "Lf": bezier_curves.Lf, "v0": bezier_curves.v0,
                                "v1": bezier_curves.v1, "D1": bezier_curves.D1}, "cases": bz}
    P5, _ = bezier_curves.get_bezier_control_points(5)
    out["bezier_sampled_100"] = {"P": P5.tolist(),
                                 "pts": [list(map(float, bezier_curves.bezier_curve(j, P5)))
                                         for j in np.linspace(0, 1, 100)]}
    # game_theory.py:92-113 (same maths, instance attributes)
    ego = game_theory.Car(name="ego", x=0, v=20, lane=1)
    other = game_theory.Car(name="fe", x=50, v=10, lane=1)
    Pg, tcag = ego.get_bezier_control_points(other, i=5)
    out["game_theory_bezier"] = {"P": Pg.tolist(), "tca": float(tcag), "ego": {"x": 0, "v": 20},
                                 "other": {"x": 50, "v": 10}, "i": 5,
                                 "car": {"h": ego.h, "W": ego.W, "theta": float(ego.θ_max), "l": ego.l, "Lf": ego.Lf}}

    # ---- dpf_test.py potentials
    dn = [{"args": [9.0, 1.4, 0, 2, 1, [10], [1.5], [0], [1]],
           "val": float(dpf_test.dnf(9.0, 1.4, 0, 2, 1, [10], [1.5], [0], [1]))}]
    for _ in range(30):
        n_obs = int(rng.integers(1, 4))
        a = [float(rng.uniform(-5, 20)), float(rng.uniform(-1, 8)), float(rng.uniform(-0.6, 0.6)),
             float(rng.uniform(0, 4)), n_obs, list(map(float, rng.uniform(0, 20, n_obs))),
             list(map(float, rng.uniform(0, 7, n_obs))), list(map(float, rng.uniform(-0.5, 0.5, n_obs))),
             list(map(float, rng.uniform(0, 3, n_obs)))]
        dn.append({"args": a, "val": float(dpf_test.dnf(*a))})
    out["dnf"] = dn
    out["u_trgl"] = [{"y": float(y), "y_t": 1.75, "val": float(dpf_test.u_trgl(y, 1.75))}
                     for y in [1.4, 0.0, 1.75, 3.3, -2.0]]
    out["u_rbd"] = [{"y": float(y), "y_br": 1, "y_bl": 6, "val": float(dpf_test.u_rbd(y, 1, 6))}
                    for y in [0.5, 1.0, 3.0, 6.0, 7.25, -1.0]]
    # event_triggered.py:26 (the module itself raises at import: line 35 calls with the wrong arity,
    # so the one-line formula is restated here from the source text)
    out["event_state"] = [{"x": 0.0, "u_x": 10.0, "a": 0.0, "t0": 1.0, "t1": 2.0,
                           "val": 0.0 + 10.0 * (1.0 + 2.0) + 1 / 2 * 0.0 * (1.0 + 2.0) ** 2},
                          {"x": 3.0, "u_x": 8.0, "a": -1.5, "t0": 0.5, "t1": 0.25,
                           "val": 3.0 + 8.0 * (0.5 + 0.25) + 1 / 2 * -1.5 * (0.5 + 0.25) ** 2}]

    with open(OUT, "w") as f:
        json.dump(out, f, indent=0)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
