"""Generate tests/golden/reference_game_theory.json by IMPORTING the reference's own game_theory.py.

Run in the build container only (needs /root/reference):   python tests/golden/gen_golden_game.py

game_theory.py:115-244: `Car.get_safety_distance`, `get_safety_payoff`, `get_velocity_payoff`, `get_total_payoff`
(which reads the module-global `ego`, game_theory.py:228), the lane decision of its __main__ loop (:377-379: change when
the lane-2 payoff exceeds the lane-1 payoff) and `get_bezier_control_points` (:92-113) towards the car in front.
Cases: the three scenario builders (:247-349) advanced along the reference's own 5 s sweep, and random scenarios around them.
"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from gen_golden import REF, _stub_matplotlib  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_game_theory.json")


def case(gt, ego, cars):
    gt.ego = ego  # the module-global that get_total_payoff reads
    rec = {"ego": [float(ego.x), float(ego.v), int(ego.lane)],
           "cars": [[float(c.x), float(c.v), int(c.lane)] for c in cars]}
    rec["payoff"] = [float(ego.get_total_payoff(cars, target_lane=1)), float(ego.get_total_payoff(cars, target_lane=2))]
    rec["safety"] = [float(ego.get_safety_payoff(cars, 1)), float(ego.get_safety_payoff(cars, 2))]
    rec["velocity"] = [float(ego.get_velocity_payoff(cars, 1)), float(ego.get_velocity_payoff(cars, 2))]
    front = ego.get_car_in_front(cars, ego.lane)
    if front is not None and ego.v != front.v:
        P, tca = ego.get_bezier_control_points(front)
        rec["P"], rec["tca"] = P.tolist(), float(tca)
    return rec


def main():
    _stub_matplotlib()
    sys.path.insert(0, REF)
    import game_theory as gt
    out = {"generator": "tests/golden/gen_golden_game.py", "numpy": np.__version__, "cases": []}
    for builder in (gt.get_cars_test_1, gt.get_cars_test_2, gt.get_cars_test_3):
        ego, cars = builder()
        for i in range(50):  # game_theory.py:355-372: dt = 0.1, 5 s
            if i % 5 == 0:
                out["cases"].append(case(gt, ego, cars))
            ego.move(0.1)
            for c in cars:
                c.move(0.1)
    rng = np.random.Generator(np.random.PCG64(424242))
    for _ in range(300):
        ego = gt.Car(name="ego", x=0.0, v=float(rng.uniform(5, 22)), lane=int(rng.integers(1, 3)))
        cars = []
        for k in range(int(rng.integers(1, 6))):
            cars.append(gt.Car(name=f"Car{k + 1}", x=float(rng.uniform(-60, 90)), v=float(rng.choice([0.0, rng.uniform(0, 25)])),
                               lane=int(rng.integers(1, 3))))
        out["cases"].append(case(gt, ego, np.array(cars)))
    # dpf_test.py:57-64 get_safe_distances (its module-level grid sweep runs at import: 0.4 s)
    import dpf_test
    sd = []
    for _ in range(40):
        Xe = [float(v) for v in rng.uniform(-20, 20, 4)]
        Xo = [float(v) for v in rng.uniform(-20, 20, 4)]
        xs, ys = dpf_test.get_safe_distances(Xe, Xo)
        sd.append({"X_e": Xe, "X_obs": Xo, "x_s": float(xs), "y_s": float(ys)})
    out["safe_distances"] = {"x_0": dpf_test.x_0, "y_0": dpf_test.y_0, "a_x_max": dpf_test.a_x_max, "a_y_max": dpf_test.a_y_max, "cases": sd}
    c0 = gt.Car()
    out["car_defaults"] = {"L": c0.L, "W": c0.W, "l": c0.l, "theta_max": float(c0.θ_max), "tlc": c0.tlc, "td": c0.td, "ti": c0.ti,
                           "tau": c0.τ, "a_max": c0.a_max, "h": c0.h, "Lf": c0.Lf}
    json.dump(out, open(OUT, "w"), indent=0)
    print("wrote", OUT, len(out["cases"]), "cases")


if __name__ == "__main__":
    main()
