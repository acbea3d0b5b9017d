"""GPU: solve parity made assertable for 100 % of instances (VERDICT r01 "next round" item 1).

The NLP is chaotic in rounding (nearest-vertex switches, line-search flips): two builds of the SAME oracle source
agree within 1e-5 on a third of the instances at the reference's eps = 1e-6 (the yardstick below).  So parity with
the oracle cannot be "U* within 1e-5 for every instance".  What CAN be asserted for every instance:

  1. GPU vs GPU, bit for bit: the amount of speculation (window width W, tile width G) and the solver's three
     shortcuts (MPCB_EXACT) change no decision — `torch.equal` on U*, statuses and iteration counts;
  2. every GPU answer is certified by the ORACLE's own evaluator and solver: feasible, the same psi as the oracle
     computes at that U*, and the oracle's PANOC warm-started there can neither improve psi by more than rounding
     nor move away;
  3. the GPU-vs-oracle distribution of |dU| is no worse than oracle-vs-oracle (two builds of one source).
"""
import os

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


def _tracking(mp, N, B, seed, flags=0, **kw):
    from mpcb200 import workloads as wl
    w = wl.tracking_batch(B, N, 100, seed=seed)
    prob = mp.BatchedMpc(N=N, S=100, road_mode=mp.ROAD_TABLE, n_roads=2, flags=flags, **kw)
    return prob, w


def _solve_tracking(mp, N, B, seed, flags=0, sp_eps=None, lam=False, **kw):
    prob, w = _tracking(mp, N, B, seed, flags, **kw)
    sp = prob.default_solver_params()
    if sp_eps is not None:
        sp.eps_target = sp_eps
    lam0 = np.zeros((B, 6 * N)) if lam else None
    U, lam_o, st = prob.solve(w["x0"], w["road"], w["U0"].copy(), lam=lam0, road_idx=w["road_idx"], sp=sp)
    return prob, w, U, lam_o, st


def _solve_lane_change(mp, N, B, seed, flags_extra=0):
    from mpcb200 import workloads as wl
    lc = wl.lane_change_batch(B, N, 100, seed=seed, n_obs=2)
    prob = mp.BatchedMpc(N=N, S=100, model_id=1, road_mode=mp.ROAD_BEZIER, flags=mp.FLAG_LANE | mp.FLAG_OBST | flags_extra,
                         n_obs=2, v_ref=20.0)
    U, _, st = prob.solve(lc["x0"], lc["ctrl"], lc["U0"].copy(), obs=lc["obs"])
    return prob, lc, U, None, st


def _assert_same_solution(a, b, what):
    import torch
    (Ua, sa), (Ub, sb) = a, b
    same_U = (Ua == Ub).all(dim=1)
    print(f"{what}: bit-identical U {same_U.double().mean().item():.4f}, statuses equal "
          f"{(sa.status == sb.status).double().mean().item():.4f}, inner counts equal "
          f"{(sa.inner_iterations == sb.inner_iterations).double().mean().item():.4f}; evaluations per solve "
          f"{(sa.evals_grad + sa.evals_value).double().mean().item():.0f} vs {(sb.evals_grad + sb.evals_value).double().mean().item():.0f}")
    assert torch.equal(sa.status, sb.status), what
    assert torch.equal(sa.inner_iterations, sb.inner_iterations), what
    assert torch.equal(sa.outer_iterations, sb.outer_iterations), what
    assert torch.equal(Ua, Ub), what
    assert torch.equal(sa.psi, sb.psi) and torch.equal(sa.residual, sb.residual), what


# ------------------------------------------------------------------ 1. speculation never changes a decision
@pytest.mark.parametrize("problem", ["pacejka_B4096_N20", "kinematic_obstacles_B512_N20"])
def test_speculation_width_is_decision_preserving(mp, problem, monkeypatch):
    """Round mode: window widths W = 7 (all trials at once), 3, 1 (one trial per round, like sequential
    back-tracking) at G = 8, and the one-lane tile G = 1, give bit-identical iterates (csrc/mpcb_solve.cuh, header)."""
    monkeypatch.setenv("MPCB_MODE", "rounds")
    out = {}
    for name, tile, w in (("G8_W7", "8", "7"), ("G8_W3", "8", "3"), ("G8_W1", "8", "1"), ("G1", "1", "1")):
        monkeypatch.setenv("MPCB_TILE", tile)
        monkeypatch.setenv("MPCB_SPECW", w)
        if problem.startswith("pacejka"):
            _, _, U, _, st = _solve_tracking(mp, 20, 4096, seed=2)
        else:
            _, _, U, _, st = _solve_lane_change(mp, 20, 512, seed=78)
        out[name] = (U, st)
    for name in ("G8_W3", "G8_W1", "G1"):
        _assert_same_solution(out["G8_W7"], out[name], f"{problem}: G8_W7 vs {name}")
    # the narrow window is what saves work: fewer evaluations for the same iterates
    ev = {k: float((v[1].evals_grad + v[1].evals_value).double().mean().item()) for k, v in out.items()}
    assert ev["G8_W1"] < 0.6 * ev["G8_W7"]


def test_device_window_policy_is_decision_preserving(mp, monkeypatch):
    """The window width picked on the device from the live-tile count (the default) gives the fixed widths' bits."""
    monkeypatch.setenv("MPCB_MODE", "rounds")
    monkeypatch.setenv("MPCB_SPECW", "7")
    _, _, Ua, _, sa = _solve_tracking(mp, 20, 8192, seed=3)
    monkeypatch.delenv("MPCB_SPECW")
    monkeypatch.setenv("MPCB_SPEC_LIVE", "1024,4096")   # all three widths occur within this batch
    _, _, Ub, _, sb = _solve_tracking(mp, 20, 8192, seed=3)
    _assert_same_solution((Ua, sa), (Ub, sb), "fixed W = 7 vs device policy")


@pytest.mark.parametrize("problem", ["pacejka_B2048_N20", "pacejka_bounded_B512_N12", "kinematic_obstacles_B512_N20"])
@pytest.mark.parametrize("mode", ["rounds", "persistent"])
def test_exact_switch_changes_nothing(mp, problem, mode, monkeypatch):
    """MPCB_FLAG_EXACT turns off the three GPU-only departures from the sequential algorithm — `ls_hopeless` (early
    rejection of a candidate that cannot pass the line-search test), `ls_settle` (descent-lemma halvings decided
    without psi(x̂+)) and the evaluation cut of a blown-up state: the iterates must not depend on them."""
    monkeypatch.setenv("MPCB_MODE", mode)
    out = []
    for extra in (0, mp.FLAG_EXACT):
        if problem.startswith("pacejka_B"):
            _, _, U, _, st = _solve_tracking(mp, 20, 2048, seed=5, flags=extra)
        elif problem.startswith("pacejka_bounded"):
            _, _, U, _, st = _solve_tracking(mp, 12, 512, seed=31, flags=mp.FLAG_BOUNDED_D | extra, lam=True,
                                             D_lower=[-np.inf] * 6, D_upper=[np.inf, np.inf, np.inf, -1.0, np.inf, np.inf])
        else:
            _, _, U, _, st = _solve_lane_change(mp, 20, 512, seed=78, flags_extra=extra)
        out.append((U, st))
    _assert_same_solution(out[0], out[1], f"{problem} [{mode}]: shortcuts on vs MPCB_FLAG_EXACT")


# ------------------------------------------------------------------ 2. every GPU answer certified by the oracle
def _orc_desc(orc, model, N, road_mode, n_roads=1, flags=0, n_obs=0, D_upper=None):
    d = orc.default_desc(model, N, 100)
    d.road_mode, d.n_roads, d.flags, d.n_obs = road_mode, n_roads, flags, n_obs
    if D_upper is not None:
        for i, v in enumerate(D_upper):
            d.D_upper[i] = v
    return d


CERT_CASES = [("N12", 12, 300, None), ("N20", 20, 300, None), ("N40", 40, 96, None), ("N20_eps1e-10", 20, 200, 1e-10)]


@pytest.mark.parametrize("name,N,B,eps", CERT_CASES, ids=[c[0] for c in CERT_CASES])
def test_oracle_certifies_every_gpu_solution(mp, orc, name, N, B, eps):
    """For EVERY instance, with the oracle's own evaluator: the GPU's U* is box-feasible, the oracle's rollout gives
    the GPU's psi at U*, and U* is a fixed point of the oracle's forward-backward map at the GPU's final step size:
    || Pi_C(U* - gamma grad psi(U*)) - U* ||_2 <= 2 eps.  (PANOC stops on ||x̂ - x|| <= eps and returns x̂; one more
    forward-backward step from x̂ is at most (1 + gamma L) times shorter-or-equal, gamma L <= 0.95.)  A full
    re-solve from U* is NOT a certificate: it re-estimates gamma and may legitimately leave for another minimiser."""
    prob, w, U, _, st = _solve_tracking(mp, N, B, seed=140 + N, sp_eps=eps)
    gam = prob.last_final_gamma(B).cpu().numpy()
    U = U.cpu().numpy()
    status = st.status.cpu().numpy()
    psi_g = st.psi.cpu().numpy()
    res_g = st.residual.cpu().numpy()
    eps_v = 1e-6 if eps is None else eps
    lb, ub = (t.cpu().numpy() for t in prob.default_box())
    assert (U >= lb).all() and (U <= ub).all()
    assert (status == 1).all(), np.bincount(status)
    assert (res_g <= eps_v).all()
    d = _orc_desc(orc, 0, N, 1, 2)
    r, psi_o = orc.fb_residual(d, w["x0"], U, w["road"], gam, road_idx=w["road_idx"])
    print(f"{name}: oracle-evaluated fixed-point residual at the GPU's (U*, gamma): r/eps median {np.median(r) / eps_v:.3f} "
          f"p99 {np.quantile(r, .99) / eps_v:.3f} max {r.max() / eps_v:.3f} (GPU's own ||p||/eps max {res_g.max() / eps_v:.3f}); "
          f"gamma median {np.median(gam):.2e} min {gam.min():.2e}; psi rel diff max {np.abs(psi_o - psi_g).max() / np.abs(psi_g).max():.1e}")
    np.testing.assert_allclose(psi_g, psi_o, rtol=1e-9)
    assert np.isfinite(gam).all() and (gam > 0).all()
    # rounding floor: the two gradients agree to ~1e-9 of their norm (tests/test_gpu_parity.py), times gamma
    assert (r <= 2.0 * eps_v + 1e-9).all(), f"worst residual {r.max():.3e} at instance {int(np.argmax(r))}"


def test_bounded_D_solve_meets_the_oracle(mp, orc):
    """ALM with a bounded D (main.py:46-51 state constraints, here vx^2 - 2 <= -1): GPU cold solve against the oracle's
    cold solve of the same problem — statuses, constraint satisfaction certified by the oracle's rollout of the GPU's
    U*, multipliers and U* within the spread the unbounded problem shows (VERDICT r01 weak point 3)."""
    N, B = 12, 200
    Dup = [np.inf, np.inf, np.inf, -1.0, np.inf, np.inf]
    prob, w, U, lam, st = _solve_tracking(mp, N, B, seed=31, flags=mp.FLAG_BOUNDED_D, lam=True, D_lower=[-np.inf] * 6,
                                          D_upper=Dup)
    U, lam = U.cpu().numpy(), lam.cpu().numpy()
    d = _orc_desc(orc, 0, N, 1, 2, flags=mp.FLAG_BOUNDED_D, D_upper=Dup)
    r = orc.solve(d, orc.default_solver_params(N), w["x0"], w["road"], w["U0"], lam=np.zeros((B, 6 * N)), road_idx=w["road_idx"])
    sg = st.status.cpu().numpy()
    # instances that start faster than the bound allows are infeasible (the car cannot brake in time): both sides must
    # give up on the same ones, with the same status
    agree = sg == r["status"]
    ok = (sg == 1) & (r["status"] == 1)
    print(f"bounded D: statuses GPU {np.bincount(sg)} oracle {np.bincount(r['status'])}, equal on {agree.mean():.3f}")
    assert agree.mean() >= 0.97 and ok.mean() > 0.5
    # constraint satisfaction of the GPU's answer, by the oracle's rollout: vx^2 - 2 <= -1 + delta at every stage
    states = orc.rollout(_orc_desc(orc, 0, N, 1, 2), w["x0"], U, w["road"], road_idx=w["road_idx"])[0]
    viol = np.maximum(states[:, :, 3] ** 2 - 2.0 + 1.0, 0.0).max(axis=1)[ok]
    active = (lam.reshape(B, N, 6)[:, :, 3] > 0).any(axis=1)[ok]
    err = np.abs(U - r["U"]).max(axis=1)[ok]
    lerr = (np.abs(lam - r["lam"]).max(axis=1) / (1.0 + np.abs(r["lam"]).max(axis=1)))[ok]
    go, gi = st.outer_iterations.cpu().numpy(), st.inner_iterations.cpu().numpy()
    same = ((go == r["outer"]) & (gi == r["inner"]))[ok]
    print(f"bounded D (both converged, {ok.sum()} instances): constraint active in {active.mean():.2f}; violation max {viol.max():.2e} "
          f"(delta 1e-4); |dU| median {np.median(err):.1e} p95 {np.quantile(err, .95):.1e}; multiplier rel err median "
          f"{np.median(lerr):.1e} p95 {np.quantile(lerr, .95):.1e}; identical outer+inner counts {same.mean():.3f}")
    assert active.mean() >= 0.05                     # the case exercises the constraint (most feasible starts are below the bound)
    assert (viol <= 1e-4 * (1 + 1e-9)).all()         # ALM's delta (controller.py:42)
    assert np.median(err) < 3e-4 and np.quantile(err, 0.95) < 5e-3   # the bars of the unbounded solve at eps = 1e-6
    assert np.median(lerr) < 1e-3
    assert abs(np.median(go[ok]) - np.median(r["outer"][ok])) <= 1


# ------------------------------------------------------------------ 3. the yardstick: oracle vs oracle
@pytest.mark.parametrize("N,B", [(12, 256), (20, 256)])
def test_gpu_vs_oracle_no_worse_than_oracle_vs_oracle(mp, orc, N, B):
    """Two builds of ONE oracle source (oracle/Makefile: -O2 without FMA contraction vs -O3 with AVX2 FMAs) on the same
    seeds give the spread that rounding alone causes on this NLP; the GPU must sit inside it (factor 2 + a floor)."""
    prob, w, U, _, st = _solve_tracking(mp, N, B, seed=40 + N)
    d = _orc_desc(orc, 0, N, 1, 2)
    sp = orc.default_solver_params(N)
    a = orc.solve(d, sp, w["x0"], w["road"], w["U0"], road_idx=w["road_idx"])
    with orc.variant("fast"):
        b = orc.solve(d, sp, w["x0"], w["road"], w["U0"], road_idx=w["road_idx"])
    U = U.cpu().numpy()
    e_oo = np.abs(a["U"] - b["U"]).max(axis=1)
    e_go = np.abs(U - a["U"]).max(axis=1)
    gi = st.inner_iterations.cpu().numpy()
    same_oo, same_go = np.mean(a["inner"] == b["inner"]), np.mean(gi == a["inner"])
    q = [0.5, 0.9, 0.99]
    print(f"N={N}: oracle vs oracle |dU| q50/90/99 {np.quantile(e_oo, q)}, within 1e-5 {np.mean(e_oo < 1e-5):.3f}, same inner "
          f"{same_oo:.3f};  GPU vs oracle {np.quantile(e_go, q)}, within 1e-5 {np.mean(e_go < 1e-5):.3f}, same inner {same_go:.3f}")
    for qq in q:
        assert np.quantile(e_go, qq) <= 2.0 * np.quantile(e_oo, qq) + 2e-5
    assert np.mean(e_go < 1e-5) >= np.mean(e_oo < 1e-5) - 0.15
    assert same_go >= same_oo - 0.15
    assert (st.status.cpu().numpy() == 1).all() and (a["status"] == 1).all()


# ------------------------------------------------------------------ 4. bounds and race evidence without compute-sanitizer
@pytest.mark.parametrize("mode", ["rounds", "persistent"])
def test_guard_canaries_and_repeat_determinism(mp, mode, monkeypatch):
    """compute-sanitizer is closed on the B200 pool (its runs left GPUs needing a reset), so: (a) every scratch buffer of
    a handle created under MPCB_GUARDS=1 sits between canary regions that must still hold their pattern after solves of
    every kind (no kernel writes outside its buffers); (b) five runs of the same batch, interleaved with other work on
    the handle, return identical bits — tiles share workspace rows and publish results to each other through global
    memory behind tile barriers, so a missing barrier or a racy read would show up as run-to-run differences."""
    import torch
    from mpcb200 import workloads as wl
    monkeypatch.setenv("MPCB_GUARDS", "1")
    monkeypatch.setenv("MPCB_MODE", mode)
    N, B = 20, 777   # ragged
    w = wl.tracking_batch(B, N, 100, seed=99)
    prob = mp.BatchedMpc(N=N, S=100, road_mode=mp.ROAD_TABLE, n_roads=2)
    outs = []
    for rep in range(5):
        U, _, st = prob.solve(w["x0"], w["road"], w["U0"].copy(), road_idx=w["road_idx"])
        outs.append((U.clone(), st.status.clone(), st.inner_iterations.clone(), st.psi.clone()))
        # other shapes through the same handle in between: smaller batch (stale workspace contents), evaluation calls
        prob.solve(w["x0"][:100 + rep], w["road"], w["U0"][:100 + rep].copy(), road_idx=w["road_idx"][:100 + rep])
        prob.value_grad(w["x0"], w["U0"], w["road"], road_idx=w["road_idx"])
        prob.rollout(w["x0"], w["U0"], w["road"], road_idx=w["road_idx"])
    for o in outs[1:]:
        for a, b in zip(outs[0], o):
            assert torch.equal(a, b)
    assert prob.check_guards() == 0
    # the other problem families on guarded handles
    bd = mp.BatchedMpc(N=12, S=100, road_mode=mp.ROAD_TABLE, n_roads=2, flags=mp.FLAG_BOUNDED_D, D_lower=[-np.inf] * 6,
                       D_upper=[np.inf, np.inf, np.inf, -1.0, np.inf, np.inf])
    w12 = wl.tracking_batch(200, 12, 100, seed=31)
    bd.solve(w12["x0"], w12["road"], w12["U0"].copy(), lam=np.zeros((200, 72)), road_idx=w12["road_idx"])
    assert bd.check_guards() == 0
    lc = wl.lane_change_batch(300, 40, 100, seed=5, n_obs=2)
    kin = mp.BatchedMpc(N=40, S=100, model_id=1, road_mode=mp.ROAD_BEZIER, flags=mp.FLAG_LANE | mp.FLAG_OBST, n_obs=2, v_ref=20.0)
    kin.solve(lc["x0"], lc["ctrl"], lc["U0"].copy(), obs=lc["obs"])
    assert kin.check_guards() == 0


# ------------------------------------------------------------------ 5. the parallel-in-time evaluation of thin rounds
@pytest.mark.parametrize("problem", ["pacejka_B4_N20", "pacejka_B40_N20", "pacejka_B2048_N12", "pacejka_bounded_B6_N12",
                                     "pacejka_B12_N40", "kinematic_obstacles_B8_N40"])
def test_parallel_in_time_evaluation_changes_nothing(mp, problem, monkeypatch):
    """Thin rounds give every gradient request a CTA of its own: the primal rollout by one thread, then all stage
    Jacobians and stage costs at once from shared memory, then the adjoint (eval_fm_pit); with fewer requests than SMs
    the psi-only requests go the same way (MPCB_PIT=2: gradient requests only).  Same stage arithmetic, same summation
    orders: bit-identical to the chunked evaluations (MPCB_PIT=0) — for a lone vehicle (every round thin), for batches
    whose tail is thin, with ALM rows (psi + states requests), and for N = 40 (two passes over the stage groups)."""
    monkeypatch.setenv("MPCB_MODE", "rounds")
    out = []
    for pit in ("0", "2", "1"):
        monkeypatch.setenv("MPCB_PIT", pit)
        if problem == "pacejka_B4_N20":
            _, _, U, _, st = _solve_tracking(mp, 20, 4, seed=8)
        elif problem == "pacejka_B40_N20":
            _, _, U, _, st = _solve_tracking(mp, 20, 40, seed=9)
        elif problem == "pacejka_B2048_N12":
            _, _, U, _, st = _solve_tracking(mp, 12, 2048, seed=10)
        elif problem == "pacejka_B12_N40":
            _, _, U, _, st = _solve_tracking(mp, 40, 12, seed=12)
        elif problem == "pacejka_bounded_B6_N12":
            _, _, U, _, st = _solve_tracking(mp, 12, 6, seed=31, flags=mp.FLAG_BOUNDED_D, lam=True, D_lower=[-np.inf] * 6,
                                             D_upper=[np.inf, np.inf, np.inf, -1.0, np.inf, np.inf])
        else:
            _, _, U, _, st = _solve_lane_change(mp, 40, 8, seed=78)
        out.append((U, st))
    _assert_same_solution(out[0], out[1], f"{problem}: chunked evaluations vs parallel-in-time gradient requests")
    _assert_same_solution(out[0], out[2], f"{problem}: chunked evaluations vs parallel-in-time for every request")


# ------------------------------------------------------------------ 6. the device-driven round loop (CUDA graph WHILE node)
@pytest.mark.parametrize("problem", ["pacejka_B1_N12", "pacejka_B3000_N20", "pacejka_bounded_B200_N12", "kinematic_obstacles_B300_N20"])
def test_graph_driven_loop_equals_host_loop(mp, problem, monkeypatch):
    """MPCB_LOOP=graph: one graph launch per solve — WHILE(any request) { transition; evaluation; advance } with the round
    index, the stop test and the choice of the transition grid on the device.  Same kernels on the same data: the same bits
    as the host-driven loop; and the call is asynchronous (it returns while the GPU is still solving)."""
    import time
    import torch
    monkeypatch.setenv("MPCB_MODE", "rounds")
    out, t_call, t_total = [], {}, {}
    for loop in ("host", "graph"):
        monkeypatch.setenv("MPCB_LOOP", loop)
        if problem == "pacejka_B1_N12":
            args = (12, 1, 8)
        elif problem == "pacejka_B3000_N20":
            args = (20, 3000, 9)
        if problem.startswith("pacejka_B"):
            prob, w = _tracking(mp, args[0], args[1], seed=args[2])
            dev = prob.device
            x0, road = torch.tensor(w["x0"], device=dev), torch.tensor(w["road"], device=dev)
            ridx = torch.tensor(w["road_idx"], device=dev)
            prob.solve(x0, road, torch.tensor(w["U0"], device=dev), road_idx=ridx)   # warm: allocations, graph build paths
            U = torch.tensor(w["U0"], device=dev)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            U, _, st = prob.solve(x0, road, U, road_idx=ridx)
            t_call[loop] = time.perf_counter() - t0
            torch.cuda.synchronize()
            t_total[loop] = time.perf_counter() - t0
            prof = prob.last_solve_profile()
            assert prof["mode"] == "rounds" and prof["rounds"] > 0 and prof["eval_ms"] > 0
        elif problem.startswith("pacejka_bounded"):
            _, _, U, _, st = _solve_tracking(mp, 12, 200, seed=31, flags=mp.FLAG_BOUNDED_D, lam=True, D_lower=[-np.inf] * 6,
                                             D_upper=[np.inf, np.inf, np.inf, -1.0, np.inf, np.inf])
        else:
            _, _, U, _, st = _solve_lane_change(mp, 20, 300, seed=78)
        out.append((U, st))
    _assert_same_solution(out[0], out[1], f"{problem}: host-driven vs graph-driven rounds")
    if problem == "pacejka_B3000_N20":
        print(f"call returned after {t_call['graph'] * 1e3:.1f} ms of a {t_total['graph'] * 1e3:.1f} ms solve (host loop: "
              f"{t_call['host'] * 1e3:.1f} of {t_total['host'] * 1e3:.1f} ms)")
        assert t_call["graph"] < 0.25 * t_total["graph"]       # asynchronous on the caller's stream
        assert t_call["host"] > 0.9 * t_total["host"]          # the host loop blocks


# ------------------------------------------------------------------ 7. safe-distance ellipse constraints
def _ellipse_problem(mp, orc, B, N, seed, n_obs=1):
    """Kinematic-4 lane change on a Bezier road past a slower car, the obstacle as the `get_safe_distances` ellipse
    CONSTRAINT (dpf_test.py:57-64, MPCB_FLAG_SAFE_DIST) instead of the `dnf` potential."""
    from mpcb200 import workloads as wl
    lc = wl.lane_change_batch(B, N, 100, seed=seed, n_obs=n_obs)
    flags = mp.FLAG_LANE | mp.FLAG_SAFE_DIST
    prob = mp.BatchedMpc(N=N, S=100, model_id=1, road_mode=mp.ROAD_BEZIER, flags=flags, n_obs=n_obs, v_ref=20.0)
    d = orc.default_desc(1, N, 100)
    d.road_mode, d.flags, d.n_obs, d.v_ref = 3, flags, n_obs, 20.0
    return prob, d, lc


@pytest.mark.parametrize("model", [1, 0])
def test_ellipse_rows_value_and_gradient_parity(mp, orc, model):
    from mpcb200 import workloads as wl
    B, N, n_obs = 64, 20, 2
    rng = np.random.default_rng(5)
    if model == 1:
        prob, d, lc = _ellipse_problem(mp, orc, B, N, seed=7, n_obs=n_obs)
        x0, road, obs, ridx = lc["x0"], lc["ctrl"], lc["obs"], None
    else:
        w = wl.tracking_batch(B, N, 100, seed=7)
        flags = mp.FLAG_SAFE_DIST | mp.FLAG_BOUNDED_D
        Dup = [np.inf, np.inf, np.inf, -1.0, np.inf, np.inf]
        prob = mp.BatchedMpc(N=N, S=100, road_mode=mp.ROAD_TABLE, n_roads=2, flags=flags, n_obs=n_obs, D_upper=Dup)
        d = _orc_desc(orc, 0, N, 1, 2, flags=flags, n_obs=n_obs, D_upper=Dup)
        x0, road, ridx = w["x0"], w["road"], w["road_idx"]
        obs = np.zeros((B, n_obs, 5))
        obs[:, :, 0] = x0[:, None, 0] + rng.uniform(0.3, 1.5, (B, n_obs))
        obs[:, :, 1] = x0[:, None, 1] + rng.uniform(-0.3, 0.3, (B, n_obs))
        obs[:, :, 3] = rng.uniform(0.2, 0.8, (B, n_obs))
    R = prob.rows
    assert R == prob.nx + n_obs and prob.m == R * N
    U = np.zeros((B, 2 * N))
    U[:, 0::2], U[:, 1::2] = rng.uniform(0, 1, (B, N)), rng.uniform(-0.1, 0.1, (B, N))
    lam, Sig = rng.uniform(0, 2, (B, R * N)), np.full((B, R * N), 10.0)
    c, g = prob.value_grad(x0, U, road, road_idx=ridx, obs=obs, lam=lam, Sigma=Sig)
    c_r, g_r = orc.value_grad(d, x0, U, road, road_idx=ridx, obs=obs, lam=lam, Sigma=Sig)
    np.testing.assert_allclose(c.cpu().numpy(), c_r, rtol=1e-9)
    scale = np.abs(g_r).max(axis=1, keepdims=True)
    assert (np.abs(g.cpu().numpy() - g_r) / scale).max() < 1e-7


def test_ellipse_constrained_solve_meets_the_oracle(mp, orc, monkeypatch):
    """ALM solve with the ellipse rows against the oracle's cold solve: same statuses, the GPU's answers satisfy the
    constraint to the ALM's delta by the oracle's own rollout, multipliers and U* within the spread of the other solves;
    and the answer does not depend on the loop (graph / host), the execution mode's speculation width or the EXACT switch."""
    import torch
    B, N = 96, 20
    prob, d, lc = _ellipse_problem(mp, orc, B, N, seed=11)
    R = prob.rows
    # put every obstacle where the uncontrolled rollout (U0) ends, in the ego's lane and slower, so that its ellipse
    # overlaps the end of the horizon and the solver has to brake or swerve
    end = orc.rollout(d, lc["x0"], lc["U0"], lc["ctrl"], obs=lc["obs"])[0][:, -1]
    rng = np.random.default_rng(11)
    v_obs = end[:, 3] * rng.uniform(0.5, 0.9, B)
    xs_end = 2.5 + (end[:, 3] - v_obs) ** 2 / 6.0
    lc["obs"][:, 0, 0] = end[:, 0] + rng.uniform(0.5, 0.9, B) * xs_end - v_obs * N * 0.05   # position at t = 0
    lc["obs"][:, 0, 1] = end[:, 1] + rng.uniform(-0.3, 0.3, B)
    lc["obs"][:, 0, 2] = 0.0
    lc["obs"][:, 0, 3] = v_obs
    lc["obs"][:, 0, 4] = 0.0
    U, lam, st = prob.solve(lc["x0"], lc["ctrl"], lc["U0"].copy(), lam=np.zeros((B, R * N)), obs=lc["obs"])
    r = orc.solve(d, orc.default_solver_params(N), lc["x0"], lc["ctrl"], lc["U0"], lam=np.zeros((B, R * N)), obs=lc["obs"])
    sg = st.status.cpu().numpy()
    agree = sg == r["status"]
    ok = (sg == 1) & (r["status"] == 1)
    Un, lamn = U.cpu().numpy(), lam.cpu().numpy()
    # constraint satisfaction of the GPU's answers by the oracle: err_z of the ellipse rows <= delta
    states = orc.rollout(d, lc["x0"], Un, lc["ctrl"], obs=lc["obs"])[0]
    t = (np.arange(N) + 1) * 0.05
    ob = lc["obs"][:, 0]
    xo = ob[:, None, 0] + (ob[:, None, 3] * t + 0.5 * ob[:, None, 4] * t * t) * np.cos(ob[:, None, 2])
    yo = ob[:, None, 1] + (ob[:, None, 3] * t + 0.5 * ob[:, None, 4] * t * t) * np.sin(ob[:, None, 2])
    vo = ob[:, None, 3] + ob[:, None, 4] * t
    vx, vy = states[:, :, 3] * np.cos(states[:, :, 2]), states[:, :, 3] * np.sin(states[:, :, 2])
    xs = 2.5 + (vx - vo * np.cos(ob[:, None, 2])) ** 2 / 6.0
    ys = 1.5 + (vy - vo * np.sin(ob[:, None, 2])) ** 2 / 2.0
    gval = 1.0 - ((states[:, :, 0] - xo) / xs) ** 2 - ((states[:, :, 1] - yo) / ys) ** 2
    viol = np.maximum(gval, 0.0).max(axis=1)
    active = (lamn.reshape(B, N, R)[:, :, prob.nx:] > 0).any(axis=(1, 2))
    err = np.abs(Un - r["U"]).max(axis=1)
    print(f"ellipse rows: statuses GPU {np.bincount(sg)} oracle {np.bincount(r['status'])}, equal {agree.mean():.3f}; constraint active "
          f"in {active[ok].mean():.2f}; violation max {viol[ok].max():.2e}; |dU| median {np.median(err[ok]):.1e} p95 {np.quantile(err[ok], .95):.1e}")
    gap = np.abs(st.psi.cpu().numpy() - r["psi"])[ok] / (1.0 + np.abs(r["psi"][ok]))
    lerr = (np.abs(lamn - r["lam"]).max(axis=1) / (1.0 + np.abs(r["lam"]).max(axis=1)))[ok]
    print(f"   psi rel gap median {np.median(gap):.1e} p95 {np.quantile(gap, .95):.1e}; multiplier rel err median {np.median(lerr):.1e}; "
          f"outer iterations GPU median {np.median(st.outer_iterations.cpu().numpy()[ok]):.0f} oracle {np.median(r['outer'][ok]):.0f}")
    assert agree.mean() >= 0.9 and ok.mean() > 0.5
    assert active[ok].mean() > 0.2                          # the constraint is exercised
    assert (viol[ok] <= 1e-4 * (1 + 1e-6)).all()            # feasible to the ALM's delta, by the oracle's rollout
    assert np.median(gap) < 1e-4                            # the same optimum value
    # bit-equality across the loop, the window width and the EXACT switch for this problem class
    ref = (U, st)
    for env in ({"MPCB_LOOP": "host"}, {"MPCB_SPECW": "1"}, {"MPCB_EXACT": "1"}, {"MPCB_MODE": "persistent", "MPCB_EXACT": "1"}):
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        p2, _, _ = _ellipse_problem(mp, orc, B, N, seed=11)
        U2, lam2, st2 = p2.solve(lc["x0"], lc["ctrl"], lc["U0"].copy(), lam=np.zeros((B, R * N)), obs=lc["obs"])
        for k in env:
            monkeypatch.delenv(k)
        if env.get("MPCB_MODE") == "persistent":   # the two modes use different psi-only kernels: compare to tolerance
            same = (st2.inner_iterations == st.inner_iterations).cpu().numpy()
            assert same.mean() > 0.9 and float((U2 - U).abs().max(dim=1).values.cpu().numpy()[same].max()) < 1e-6
        else:
            _assert_same_solution(ref, (U2, st2), f"ellipse rows: default vs {env}")
            assert torch.equal(lam, lam2)
