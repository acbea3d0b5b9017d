"""GPU: the reference's operator API itself — `MPCController`, `create_casadi_problem`,
`alpaqa_vehicle_test`, the alpaqa-shaped solver objects, the legacy controllers and the reference-named model / road
façades — exercised the way reference main.py / controller.py use them, checked against the oracle and the
reference-generated golden vectors."""
import warnings

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mp():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("these tests need a CUDA device (there is no CPU fallback to test instead)")
    import mpcb200
    return mpcb200


def test_alpaqa_vehicle_test_matches_oracle_driven_loop(mp, orc):
    """reference main.py:62-167 through the drop-in modules (20 ticks, N_horiz = 12, straight road, the reference's
    start) against the same loop driven by the oracle: solve from the previous (U, lambda) unshifted
    (controller.py:57), apply the first input, one RK4 stage of the plant (main.py:145)."""
    from mpcb200 import main as ref_main
    N_sim, N = 20, 12
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        y_mpc, ctl = ref_main.alpaqa_vehicle_test(N_sim=N_sim, N_horiz=N)
    assert y_mpc.shape == (6, N_sim)
    assert ctl.failures == 0 and ctl.tot_it > 0
    assert ctl.U.shape == (2 * N,) and ctl.λ.shape == (6 * N,)
    assert ctl.N_horiz == N and ctl.u_dim == 2 and ctl.solver is not None and ctl.problem is not None
    # oracle-driven loop
    d = orc.default_desc(0, N, 100)
    sp = orc.default_solver_params(N)
    road = ref_main.get_centerline(100, True).ravel(order="F")
    y = np.array([0, 0, 0, .5, 0, 0], dtype=float)
    U = np.tile([1.0, 0.0], N)
    ref = np.empty((6, N_sim))
    tot = 0
    for n in range(N_sim):
        r = orc.solve(d, sp, y[None], road, U[None])
        assert r["status"][0] == 1
        U = r["U"][0]
        tot += int(r["inner"][0])
        y = orc.rk4_stage(d, y, U[:2])
        ref[:, n] = y
    err = np.abs(y_mpc - ref).max()
    print(f"closed loop, 20 ticks: max |state - oracle loop| = {err:.2e}; inner iterations {ctl.tot_it} vs oracle {tot}")
    assert err < 1e-6
    assert abs(ctl.tot_it - tot) <= 0.1 * tot + 2


def test_batched_controller_call_equals_batched_solve(mp):
    """`MPCController((B, 6), (2S,))` (the batched extension of controller.py:51) returns what BatchedMpc.solve returns
    for the same inputs, takes the SHARED road layout for one centerline and a road TABLE for a few, and keeps the
    reference's attributes (U, λ, tot_it, failures: controller.py:18-21,63-64)."""
    from mpcb200 import main as ref_main, workloads as wl
    from mpcb200.car_dynamics import KinematicBicyclePacejka
    from mpcb200.controller import MPCController
    N, B, S = 12, 64, 100
    model = KinematicBicyclePacejka()
    w = wl.tracking_batch(B, N, S, seed=17)
    straight = w["road"][0]
    # (1) one shared centerline
    prob = ref_main.create_casadi_problem(model, N, S, 1.0, 1.0, 0.32)
    prob.param = np.concatenate((w["x0"][0], straight, ref_main.VEHICLE_PARAM))
    ctl = MPCController(model, prob, N)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        U = ctl(w["x0"], straight)
    assert U.shape == (B, 2 * N) and ctl.U.shape == (B, 2 * N) and ctl.λ.shape == (B, 6 * N)
    assert ctl.last_stats["road_layout"] == "shared"
    ref = mp.BatchedMpc(N=N, S=S)
    Ur, _, st = ref.solve(w["x0"], straight, w["U0"].copy())
    assert (U == Ur.cpu().numpy()).all()
    assert ctl.tot_it == int(st.inner_iterations.sum().item())
    assert ctl.failures == int((st.status != 1).sum().item())
    assert (ctl.last_stats["inner_convergence_failures"] >= 0).all()
    # (2) per-instance centerlines drawn from two roads: deduplicated into a table
    prob2 = ref_main.create_casadi_problem(model, N, S, 1.0, 1.0, 0.32)
    prob2.param = np.concatenate((w["x0"][0], straight, ref_main.VEHICLE_PARAM))
    ctl2 = MPCController(model, prob2, N)
    per = w["road"][w["road_idx"]]
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        U2 = ctl2(w["x0"], per)
    assert ctl2.last_stats["road_layout"] == "table"
    ref2 = mp.BatchedMpc(N=N, S=S, road_mode=mp.ROAD_TABLE, n_roads=2)
    Ur2, _, _ = ref2.solve(w["x0"], w["road"], w["U0"].copy(), road_idx=w["road_idx"])
    assert (U2 == Ur2.cpu().numpy()).all()
    # (3) second tick: warm start from the previous solution, unshifted (controller.py:57)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        U3 = ctl(w["x0"], straight)
    Ur3, _, st3 = ref.solve(w["x0"], straight, Ur.clone())
    assert (U3 == Ur3.cpu().numpy()).all()
    assert np.median(st3.inner_iterations.cpu().numpy()) <= np.median(st.inner_iterations.cpu().numpy())


def test_progress_callback_and_heuristic_warning(mp):
    """controller.py:23-24,38: the per-iteration callback receives psi of every inner iteration (replayed from the
    kernel's trace after the solve); controller.py:32's hessian_step_size_heuristic is ignored with a warning."""
    import mpcb200.alpaqa_compat as pa
    from mpcb200 import main as ref_main
    from mpcb200.car_dynamics import KinematicBicyclePacejka
    N, S = 12, 100
    model = KinematicBicyclePacejka()
    prob = ref_main.create_casadi_problem(model, N, S, 1.0, 1.0, 0.32)
    road = ref_main.get_centerline(S, True).ravel(order="F")
    prob.param = np.concatenate(([0, .3, .2, .8, 0, 0], road, ref_main.VEHICLE_PARAM))
    pa._warned_heuristic = False
    inner = pa.StructuredPANOCLBFGSSolver(panoc_params={"stop_crit": pa.ProjGradNorm2, "max_iter": 1000,
                                                        "hessian_step_size_heuristic": 15}, lbfgs_params={"memory": N})
    seen = []
    inner.set_progress_callback(lambda info: seen.append((info.k, info.ψ)))
    solver = pa.ALMSolver(alm_params={"ε": 1e-6, "δ": 1e-4, "Σ_0": 1e5, "max_iter": 1000}, inner_solver=inner)
    with pytest.warns(UserWarning, match="hessian_step_size_heuristic"):
        x, y, stats = solver(prob, np.tile([1.0, 0.0], N), np.zeros(6 * N))
    assert stats["status"] == pa.SolverStatus.Converged
    psis = np.array([p for _, p in seen])
    print(f"callback fired {len(seen)} times for {stats['inner']['iterations']} inner iterations; psi {psis[0]:.4f} -> {psis[-1]:.6f}")
    assert len(seen) >= stats["inner"]["iterations"]
    assert np.isfinite(psis).all() and psis[-1] <= psis[0]
    assert psis[-1] == pytest.approx(stats["ψ"], rel=1e-6)
    assert stats["inner_convergence_failures"] == 0
    assert "hessian_step_size_heuristic" in stats


def test_legacy_controllers_are_importable_and_run(mp):
    """controller.py:72-135 (`mpc_controller`, `mpc_controller_initial`) and :138 (`straight_line_controller`) with the
    reference-named model and road objects."""
    from mpcb200.controller import mpc_controller, mpc_controller_initial, straight_line_controller
    from mpcb200.dynamics import KinematicBicyclePacejka
    from mpcb200.road import Road
    model, road = KinematicBicyclePacejka(), Road()
    x = np.array([5.0, 5.0, np.pi / 2, 0.8, 0.0, 0.0])
    u = mpc_controller(model, x, road, 1.0, N=2, dt=0.1)
    assert u.shape == (2,) and np.isfinite(u).all()
    u0 = mpc_controller_initial(model, x, x + np.array([0.1, 0, 0, 0, 0, 0]), N=2, dt=0.1)
    assert u0.shape == (2,) and np.isfinite(u0).all()
    assert (straight_line_controller(x, road) == np.array([1, 0])).all()


def test_reference_named_facades_match_golden(mp, golden):
    """dynamics.py / road.py / car_dynamics.py façades (each call launches a kernel) against the vectors the reference's
    own NumPy code produced (tests/golden/gen_golden.py)."""
    from mpcb200 import dynamics, road, car_dynamics
    for key, cls in (("pacejka", dynamics.KinematicBicyclePacejka), ("kinematic", dynamics.KinematicBicycleSimplified)):
        m = cls()
        for c in golden[key]:
            np.testing.assert_allclose(m(np.array(c["x"]), np.array(c["u"])), c["xdot"], rtol=1e-12, atol=1e-12)
    r = road.Road()
    for c in golden["road_circle"]["cases"]:
        idx, pt = r.find_nearest_point(np.array(c["pos"]))
        assert idx == c["idx"]
        cte, he, pe = r.compute_errors(np.array(c["pos"]), c["heading"])
        np.testing.assert_allclose([cte, he, pe], [c["cte"], c["heading_err"], c["pos_err"]], rtol=1e-11, atol=1e-13)
    for c in golden["wrap_to_pi"]:
        assert r.wrap_to_pi(c["a"]) == pytest.approx(c["w"], abs=1e-15)
    m = car_dynamics.KinematicBicyclePacejka()
    f_d = m.dynamics(0.05)
    y1 = f_d(np.array([0, 0, 0, .5, 0, 0.]), np.array([1.0, 0.0]))[0]
    np.testing.assert_allclose(y1[[0, 3]], [0.025808, 0.532159], atol=2e-6)   # SURVEY Appendix B RK4 anchor
    U = np.tile([1.0, 0.0], 20)
    assert m.input_to_matrix(U).shape == (2, 20)
    st = m.simulate(20, np.array([0, 0, 0, .5, 0, 0.]), U, None)
    np.testing.assert_allclose(st[[0, 3], -1], [0.771028, 0.994069], atol=2e-6)  # Appendix B: after 20 stages
