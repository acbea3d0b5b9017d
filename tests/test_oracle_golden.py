"""CPU: pin the oracle against golden vectors produced by the REFERENCE's own NumPy code
(tests/golden/gen_golden.py imported /root/reference to make them)."""
import numpy as np
import pytest

RTOL = 1e-13  # same formulas, same libm: agreement to a few ulp


def test_pacejka_ode_matches_dynamics_py(orc, golden):
    for c in golden["pacejka"]:
        out = orc.dynamics(0, c["x"], c["u"], clip=True)
        np.testing.assert_allclose(out, c["xdot"], rtol=RTOL, atol=1e-13)


def test_survey_appendix_b_pacejka(orc):
    out = orc.dynamics(0, [0, 0, 0, .5, 0, 0], [.3, .1], clip=True)
    np.testing.assert_allclose(out, [0.5, 0, 0, -0.33463410113156905, 1.1537190288651225, 51.4098460157412], rtol=1e-14)
    out = orc.dynamics(0, [0, 0, 0, .5, 0, 0], [2.0, 0.5], clip=True)  # clipped to [1, 0.32]
    np.testing.assert_allclose(out, [0.5, 0, 0, -0.5060181024462355, 3.4962245144941315, 155.79214646689834], rtol=1e-14)


def test_kinematic_ode_matches_dynamics_py(orc, golden):
    for c in golden["kinematic"]:
        out = orc.dynamics(1, c["x"], c["u"], clip=True)
        np.testing.assert_allclose(out, c["xdot"], rtol=RTOL, atol=1e-14)


@pytest.mark.parametrize("name", ["road_circle", "road_straight"])
def test_road_errors_match_road_py(orc, golden, name):
    road = np.array(golden[name]["centerline_flat_F"])
    for c in golden[name]["cases"]:
        idx, err = orc.road_errors(road, c["pos"], c["heading"], normalised=True)
        assert idx == c["idx"]
        np.testing.assert_allclose(err, [c["cte"], c["heading_err"], c["pos_err"]], rtol=1e-12, atol=1e-14)


def test_casadi_errors_are_numpy_errors_times_segment_length(orc, golden):
    # car_dynamics.py:205-227 omits the division by |w| that road.py:67,79 has; away from the ends the
    # two scans pick the same vertex, so raw = normalised * segment length.
    road = np.array(golden["road_straight"]["centerline_flat_F"])
    for c in golden["road_straight"]["cases"]:
        idx, err = orc.road_errors(road, c["pos"], c["heading"], normalised=False)
        assert idx == c["idx"]
        np.testing.assert_allclose(err[0], c["cte"] * 0.1, rtol=1e-9, atol=1e-15)
        np.testing.assert_allclose(err[2], c["pos_err"] * 0.1, rtol=1e-9, atol=1e-15)
        np.testing.assert_allclose(err[1], c["heading_err"], rtol=1e-12, atol=1e-15)


def test_nearest_scan_casadi_semantics(orc):
    # scan i = 1..S-2, strict '<', vertex 0 initial with prev = vertex 0: the LAST vertex can never win
    S = 10
    road = np.concatenate([np.arange(S, dtype=float), np.zeros(S)])
    idx, _ = orc.road_errors(road, (9.0, 0.0), 0.0, normalised=False)
    assert idx == S - 2
    idx, _ = orc.road_errors(road, (-1.0, 0.0), 0.0, normalised=False)
    assert idx == 0
    idx, err = orc.road_errors(road, (2.5, 0.3), 0.0, normalised=False)  # exact tie: first wins
    assert idx == 2
    idx_np, _ = orc.road_errors(road, (9.0, 0.0), 0.0, normalised=True)
    assert idx_np == S - 1


def test_wrap_to_pi(orc, golden):
    for c in golden["wrap_to_pi"]:
        assert orc.wrap_to_pi(c["a"]) == pytest.approx(c["w"], abs=1e-15)
    assert orc.wrap_to_pi(np.pi) == -np.pi and orc.wrap_to_pi(-np.pi) == -np.pi


def test_bezier(orc, golden):
    k = golden["bezier"]["consts"]
    for c in golden["bezier"]["cases"]:
        P, tca = orc.bezier_control_points(c["i"], k["h"], k["W"], k["theta"], k["l"], k["Lf"], k["v0"], k["v1"], k["D1"])
        np.testing.assert_allclose(P, c["P"], rtol=1e-14)
        assert tca == pytest.approx(c["tca"], rel=1e-14)
        for j, pt in zip(c["j"], c["pts"]):
            np.testing.assert_allclose(orc.bezier_point(np.array(c["P"]), j), pt, rtol=1e-13, atol=1e-13)
    s = golden["bezier_sampled_100"]
    road = orc.bezier_sample(np.array(s["P"]), 100)
    np.testing.assert_allclose(np.stack([road[:100], road[100:]], 1), s["pts"], rtol=1e-13, atol=1e-12)
    g = golden["game_theory_bezier"]
    P, tca = orc.bezier_control_points(g["i"], g["car"]["h"], g["car"]["W"], g["car"]["theta"], g["car"]["l"],
                                       g["car"]["Lf"], g["ego"]["v"], g["other"]["v"], g["other"]["x"] - g["ego"]["x"])
    np.testing.assert_allclose(P, g["P"], rtol=1e-14)
    assert tca == pytest.approx(g["tca"], rel=1e-14)


def test_potentials(orc, golden):
    for c in golden["dnf"]:
        a = c["args"]
        assert orc.dnf(*a) == pytest.approx(c["val"], rel=1e-12, abs=1e-300)
    for c in golden["u_trgl"]:
        assert orc.u_trgl(c["y"], c["y_t"]) == pytest.approx(c["val"], rel=1e-15)
    for c in golden["u_rbd"]:
        assert orc.u_rbd(c["y"], c["y_br"], c["y_bl"]) == pytest.approx(c["val"], rel=1e-15)
    for c in golden["event_state"]:
        assert orc.event_state(c["x"], c["u_x"], c["a"], c["t0"], c["t1"]) == c["val"]


def test_game_theory_payoffs_match_reference(orc):
    """game_theory.py:115-244 (`get_total_payoff` for lanes 1 and 2 incl. the car-behind term and the module-global ego,
    the lane decision, `get_bezier_control_points`) against vectors produced by the reference's own code
    (tests/golden/gen_golden_game.py): 330 cases incl. SURVEY Appendix B's test-1 payoffs."""
    import json
    import os
    g = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_game_theory.json")))
    assert g["cases"][0]["payoff"] == [0.7999999999999999, 1.0365591188311774]
    M = 8
    n_ctrl = 0
    for c in g["cases"]:
        cars = np.zeros((M, 3))
        cars[:len(c["cars"])] = c["cars"]
        r = orc.game_payoff(np.array([c["ego"]]), cars[None])
        np.testing.assert_allclose(r["payoff"][0], c["payoff"], rtol=1e-14, atol=1e-15)
        assert r["decision"][0] == int(c["payoff"][1] > c["payoff"][0])
        if "P" in c:
            np.testing.assert_allclose(r["ctrl"][0].reshape(2, 6), c["P"], rtol=1e-14, atol=1e-14)
            assert r["tca"][0] == pytest.approx(c["tca"], rel=1e-14)
            n_ctrl += 1
        else:
            assert np.isnan(r["ctrl"][0]).all()
    assert n_ctrl > 100
