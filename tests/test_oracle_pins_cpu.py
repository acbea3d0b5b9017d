"""CPU: pins of the oracle's value / gradient / solver that need no alpaqa.

  * the oracle's reverse sweep against central differences of its own rollout cost;
  * SURVEY Appendix B's NLP anchors (NumPy restatement + SciPy, made while surveying): J* = 0.75206669 with 11 of
    24 inputs at a bound for the main.py start at N = 12; J* = 0.19407432 (N = 12) / 0.2439775 (N = 20) and the first
    inputs [1, -0.1833, 1, -0.1390, 1, -0.0770, 1, -0.0360] for y0 = [0, .3, .2, .8, 0, 0];
  * SciPy L-BFGS-B on the oracle's value/gradient (the reference itself uses scipy.optimize.minimize,
    controller.py:104) and the oracle's own PANOC reach the same psi and, at a tolerance that pins U*, the same U*.
The solver itself stays "parity unpinned" against alpaqa (not installable here); these pin everything around it.
"""
import numpy as np
import pytest

S = 100


def _straight():
    from mpcb200 import workloads as wl
    return wl.flat_F(wl.get_centerline(S, True))


@pytest.mark.parametrize("model,N", [(0, 12), (0, 20), (1, 20)])
def test_adjoint_matches_central_differences(orc, model, N):
    from mpcb200 import workloads as wl
    B = 48
    w = wl.tracking_batch(B, N, S, seed=7, model="pacejka" if model == 0 else "kin", random_U=True, d_min=0.2)
    d = orc.default_desc(model, N, S)
    d.road_mode, d.n_roads = 1, 2
    _, g = orc.value_grad(d, w["x0"], w["U0"], w["road"], road_idx=w["road_idx"])
    h = 1e-6
    fd = np.zeros_like(g)
    for i in range(2 * N):
        Up, Um = w["U0"].copy(), w["U0"].copy()
        Up[:, i] += h
        Um[:, i] -= h
        cp = orc.rollout(d, w["x0"], Up, w["road"], road_idx=w["road_idx"])[2]
        cm = orc.rollout(d, w["x0"], Um, w["road"], road_idx=w["road_idx"])[2]
        fd[:, i] = (cp - cm) / (2 * h)
    err = np.abs(fd - g).max(axis=1) / np.abs(g).max(axis=1)
    print(f"model {model} N={N}: |FD - adjoint| / |g|_inf median {np.median(err):.1e} max {err.max():.1e}")
    # the cost is piecewise smooth (nearest-vertex switches): a difference that straddles a switch is off by O(1);
    # away from them the truncation + rounding error of a 1e-6 central difference is ~1e-9
    assert np.median(err) < 1e-8
    assert np.quantile(err, 0.9) < 1e-6


def _scipy_solve(orc, N, y0):
    from scipy.optimize import minimize
    d = orc.default_desc(0, N, S)
    road = _straight()
    y0 = np.array([y0], dtype=float)

    def fun(u):
        c, g = orc.value_grad(d, y0, u[None], road)
        return float(c[0]), g[0]

    lb = np.tile([-1.0, -0.32], N)  # main.py:55-56
    r = minimize(fun, np.tile([1.0, 0.0], N), jac=True, method="L-BFGS-B", bounds=list(zip(lb, -lb)),
                 options=dict(maxiter=2000, ftol=1e-15, gtol=1e-10))
    return d, road, y0, lb, r


def test_appendix_b_anchor_main_py_start(orc):
    """main.py:72-79 start, N = 12: J* = 0.75206669, 11 of 24 inputs at a bound."""
    d, road, y0, lb, r = _scipy_solve(orc, 12, [0, 0, 0, .5, 0, 0])
    assert r.fun == pytest.approx(0.75206669, abs=5e-9)
    at_bound = (np.abs(r.x - lb) < 1e-9) | (np.abs(r.x + lb) < 1e-9)
    assert at_bound.sum() == 11
    o = orc.solve(d, orc.default_solver_params(12), y0, road, np.tile([1.0, 0.0], 12)[None])
    assert o["status"][0] == 1
    assert o["psi"][0] == pytest.approx(r.fun, rel=1e-12)
    assert np.abs(o["U"][0] - r.x).max() < 1e-8


@pytest.mark.parametrize("N,J,first", [(12, 0.19407432, None),
                                       (20, 0.2439775, [1, -0.1833, 1, -0.1390, 1, -0.0770, 1, -0.0360])])
def test_appendix_b_anchor_offset_start(orc, N, J, first):
    d, road, y0, lb, r = _scipy_solve(orc, N, [0, .3, .2, .8, 0, 0])
    assert r.fun == pytest.approx(J, abs=5e-8)
    if first is not None:
        np.testing.assert_allclose(r.x[:8], first, atol=6e-5)
    # the oracle's PANOC, at the reference's eps = 1e-6 and at a tolerance that pins U*
    sp = orc.default_solver_params(N)
    o = orc.solve(d, sp, y0, road, np.tile([1.0, 0.0], N)[None])
    assert o["status"][0] == 1 and o["psi"][0] == pytest.approx(r.fun, rel=1e-8)
    assert np.abs(o["U"][0] - r.x).max() < 3e-4          # eps = 1e-6 pins U* only this far (cost is flat there)
    sp.eps_target = 1e-10
    o = orc.solve(d, sp, y0, road, np.tile([1.0, 0.0], N)[None])
    assert o["status"][0] == 1 and o["psi"][0] == pytest.approx(r.fun, rel=1e-12)
    assert np.abs(o["U"][0] - r.x).max() < 1e-5          # north-star tolerance


def test_scipy_and_oracle_panoc_reach_the_same_minimum_on_a_batch(orc):
    """Algorithm-independent check on config-2 style instances: PANOC's psi is never worse than L-BFGS-B's by more
    than rounding, and most instances land in the same minimiser (the NLP has several: nearest-vertex switches)."""
    from scipy.optimize import minimize
    from mpcb200 import workloads as wl
    N, B = 12, 24
    w = wl.tracking_batch(B, N, S, seed=21)
    d = orc.default_desc(0, N, S)
    d.road_mode, d.n_roads = 1, 2
    sp = orc.default_solver_params(N)
    sp.eps_target = 1e-10
    o = orc.solve(d, sp, w["x0"], w["road"], w["U0"], road_idx=w["road_idx"])
    lb = np.tile([-1.0, -0.32], N)
    gap, dU = [], []
    for b in range(B):
        def fun(u):
            c, g = orc.value_grad(d, w["x0"][b:b + 1], u[None], w["road"], road_idx=w["road_idx"][b:b + 1])
            return float(c[0]), g[0]
        r = minimize(fun, w["U0"][b], jac=True, method="L-BFGS-B", bounds=list(zip(lb, -lb)),
                     options=dict(maxiter=3000, ftol=1e-15, gtol=1e-10))
        gap.append((o["psi"][b] - r.fun) / abs(r.fun))
        dU.append(np.abs(o["U"][b] - r.x).max())
    gap, dU = np.array(gap), np.array(dU)
    print(f"PANOC vs L-BFGS-B: psi rel gap median {np.median(np.abs(gap)):.1e} max {gap.max():.1e}; |dU| median "
          f"{np.median(dU):.1e}; same minimiser (|dU| < 1e-5) {np.mean(dU < 1e-5):.2f}")
    assert (o["status"] == 1).all()
    assert np.median(np.abs(gap)) < 1e-9
    assert np.mean(dU < 1e-5) >= 0.75


def test_safe_distances_match_reference_and_ellipse_rows_have_the_right_gradient(orc):
    """dpf_test.py:57-64 `get_safe_distances` against vectors produced by the reference's own code, and the ellipse
    constraint rows built from it (MPCB_FLAG_SAFE_DIST: g = 1 - (dx/x_s)^2 - (dy/y_s)^2, through the ALM terms of psi)
    against central differences — both vehicle models, the state rows of main.py:46-51 active next to them."""
    import json
    import os
    from mpcb200 import workloads as wl
    g = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_game_theory.json")))
    sd = g["safe_distances"]
    assert (sd["x_0"], sd["y_0"], sd["a_x_max"], sd["a_y_max"]) == (5, 3, 3, 1)
    for c in sd["cases"]:
        assert orc.safe_distances(c["X_e"], c["X_obs"]) == (c["x_s"], c["y_s"])
    for model, N in ((1, 12), (0, 12)):
        B, n_obs = 12, 2
        d = orc.default_desc(model, N, S)
        rng = np.random.default_rng(3)
        if model == 1:
            lc = wl.lane_change_batch(B, N, S, seed=7, n_obs=n_obs)
            d.road_mode, d.flags, d.n_obs, d.v_ref = 3, 2 | 16, n_obs, 20.0
            x0, road, obs, ridx = lc["x0"], lc["ctrl"], lc["obs"], None
        else:
            w = wl.tracking_batch(B, N, S, seed=7)
            d.road_mode, d.n_roads, d.flags, d.n_obs = 1, 2, 16 | 4, n_obs
            d.D_upper[3] = -1.0
            x0, road, ridx = w["x0"], w["road"], w["road_idx"]
            obs = np.zeros((B, n_obs, 5))
            obs[:, :, 0] = x0[:, None, 0] + rng.uniform(0.3, 1.5, (B, n_obs))
            obs[:, :, 1] = x0[:, None, 1] + rng.uniform(-0.3, 0.3, (B, n_obs))
            obs[:, :, 3] = rng.uniform(0.2, 0.8, (B, n_obs))
        R = orc.rows_per_stage(d)
        assert R == (4 if model == 1 else 6) + n_obs
        U = np.zeros((B, 2 * N))
        U[:, 0::2], U[:, 1::2] = rng.uniform(0, 1, (B, N)), rng.uniform(-0.1, 0.1, (B, N))
        lam, Sig = rng.uniform(0, 2, (B, R * N)), np.full((B, R * N), 10.0)
        kw = dict(road_idx=ridx, obs=obs, lam=lam, Sigma=Sig)
        _, gr = orc.value_grad(d, x0, U, road, **kw)
        h, fd = 1e-6, np.zeros_like(gr)
        for i in range(2 * N):
            Up, Um = U.copy(), U.copy()
            Up[:, i] += h
            Um[:, i] -= h
            fd[:, i] = (orc.value_grad(d, x0, Up, road, **kw)[0] - orc.value_grad(d, x0, Um, road, **kw)[0]) / (2 * h)
        err = np.abs(fd - gr).max(axis=1) / np.abs(gr).max(axis=1)
        assert np.median(err) < 1e-7 and err.max() < 1e-5, (model, err.max())
