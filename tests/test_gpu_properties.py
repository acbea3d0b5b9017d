"""GPU: property tests of the solve (hypothesis): box feasibility of every iterate PANOC can return,
projection idempotence, decrease of the forward-backward envelope over accepted steps, invariance to the batch an
instance travels in.  (Invariance to the GPU count: tools/config4_shard_check.py under NCCL.)"""
import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings, strategies as st

pytestmark = pytest.mark.gpu
SET = dict(max_examples=8, deadline=None, suppress_health_check=list(HealthCheck), derandomize=True)


@pytest.fixture(scope="module")
def mp():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("these tests need a CUDA device (there is no CPU fallback to test instead)")
    import mpcb200
    return mpcb200


@pytest.fixture(scope="module")
def prob(mp):
    return mp.BatchedMpc(N=12, S=100, road_mode=mp.ROAD_TABLE, n_roads=2)


def _batch(seed, B=96, N=12):
    from mpcb200 import workloads as wl
    return wl.tracking_batch(B, N, 100, seed=seed)


@settings(**SET)
@given(seed=st.integers(0, 10_000), k=st.integers(0, 6), lo_d=st.floats(-1.0, -0.05), hi_d=st.floats(0.05, 1.0),
       hi_s=st.floats(0.02, 0.32), spread=st.floats(0.0, 3.0))
def test_every_returned_iterate_is_box_feasible(prob, seed, k, lo_d, hi_d, hi_s, spread):
    """PANOC returns x̂_k = Pi_C(x_k - gamma grad psi(x_k)); stopped after k = 0..6 iterations (status MaxIter) or at
    convergence, from warm starts far outside the box, for random boxes: lb <= U <= ub exactly."""
    w = _batch(seed)
    B, N = w["x0"].shape[0], 12
    rng = np.random.default_rng(seed)
    lb, ub = np.tile([lo_d, -hi_s], N), np.tile([hi_d, hi_s], N)
    U0 = rng.uniform(-1 - spread, 1 + spread, (B, 2 * N))      # mostly infeasible warm starts
    sp = prob.default_solver_params()
    sp.max_iter = k
    U, _, stt = prob.solve(w["x0"], w["road"], U0, road_idx=w["road_idx"], lb=lb, ub=ub, sp=sp)
    U = U.cpu().numpy()
    s = stt.status.cpu().numpy()
    assert np.isin(s, [1, 3, 5]).all(), np.bincount(s)         # Converged / MaxIter / NoProgress
    assert (U >= lb).all() and (U <= ub).all()


@settings(**SET)
@given(seed=st.integers(0, 10_000), spread=st.floats(0.0, 5.0))
def test_projection_is_idempotent(mp, seed, spread):
    """With every cost weight zero psi == 0, so the solve IS the projection: U* = Pi_C(U0) (clamp, exactly), and solving
    again from U* returns U* (Pi_C Pi_C = Pi_C) in zero iterations."""
    N = 12
    p0 = mp.BatchedMpc(N=N, S=100, road_mode=mp.ROAD_TABLE, n_roads=2, c=[0, 0, 0, 0, 0, 0])
    w = _batch(seed, B=64)
    rng = np.random.default_rng(seed)
    U0 = rng.uniform(-1 - spread, 1 + spread, (64, 2 * N))
    lb, ub = (t.cpu().numpy() for t in p0.default_box())
    U1, _, s1 = p0.solve(w["x0"], w["road"], U0, road_idx=w["road_idx"])
    assert (s1.status.cpu().numpy() == 1).all()
    assert (U1.cpu().numpy() == np.clip(U0, lb, ub)).all()
    U2, _, s2 = p0.solve(w["x0"], w["road"], U1.clone(), road_idx=w["road_idx"])
    assert (U2 == U1).all().item() and (s2.inner_iterations == 0).all().item()


@settings(**SET)
@given(seed=st.integers(0, 10_000))
def test_fbe_decreases_over_accepted_steps(prob, seed):
    """Appendix A 3d: a step is accepted when phi+ <= phi - sigma + (1 + |phi|) 10 eps with sigma >= 0, so along one
    PANOC run the forward-backward envelope never rises by more than that slack over a step a tau candidate produced.  Read from the per-iteration trace (psi, phi_gamma, gamma) of mpcb_set_trace;
    psi(x_k) itself may rise — PANOC is not a descent method on psi."""
    w = _batch(seed, B=64)
    prob.set_trace(512)
    try:
        U, _, stt = prob.solve(w["x0"], w["road"], w["U0"].copy(), road_idx=w["road_idx"])
        tr = prob.last_trace(64).cpu().numpy()
    finally:
        prob.set_trace(0)
    assert (stt.status.cpu().numpy() == 1).all()
    checked = 0
    for b in range(64):
        t = tr[b]
        t = t[np.isfinite(t[:, 0])]
        assert len(t) >= 1
        phi, gam, cand = t[:, 1], t[:, 2], t[:, 3]
        # iterate k+1 was produced by candidate cand[k+1]: a tau step (0..7) passed the line-search test; the prox step (8)
        # is accepted unconditionally once every tau has failed (Appendix A 3d), and where the gradient overflows the
        # step-size cap its envelope may be anything (the oracle accepts it too) — only tau steps carry the guarantee.
        # A new PANOC run of the tolerance continuation (-1) re-estimates gamma.
        ok = (cand[1:] >= 0) & (cand[1:] <= 7) & (gam[1:] <= gam[:-1])
        rise = phi[1:] - phi[:-1]
        slack = (1.0 + np.abs(phi[:-1])) * 10 * 2.220446049250313e-16
        assert (rise[ok] <= slack[ok] + 1e-300).all(), (b, rise[ok].max())
        checked += int(ok.sum())
    assert checked > 64


@settings(**SET)
@given(seed=st.integers(0, 10_000), lo=st.integers(0, 60), n=st.integers(1, 36))
def test_result_does_not_depend_on_the_batch(prob, seed, lo, n):
    """An instance's answer is the same bits whether it is solved in a batch of 96, in any sub-batch, or reversed."""
    import torch
    w = _batch(seed)
    U, _, s = prob.solve(w["x0"], w["road"], w["U0"].copy(), road_idx=w["road_idx"])
    sl = slice(lo, lo + n)
    Us, _, ss = prob.solve(w["x0"][sl], w["road"], w["U0"][sl].copy(), road_idx=w["road_idx"][sl])
    assert torch.equal(U[sl], Us) and torch.equal(s.inner_iterations[sl], ss.inner_iterations)
    rev = np.arange(96)[::-1].copy()
    Ur, _, _ = prob.solve(w["x0"][rev], w["road"], w["U0"][rev].copy(), road_idx=w["road_idx"][rev])
    assert torch.equal(U[torch.as_tensor(rev.copy(), device=U.device)], Ur)
