"""Batched drivers around the MPC solve for the workloads of SURVEY.md §8(d) configs 1, 3, 4 and §8(f)-1.

Everything here is host-side sequencing of `BatchedMpc` calls (libmpcb200.so kernels) on device tensors; the
only tensor arithmetic done in torch is bookkeeping (obstacle records, perturbations, shifts).  Nothing falls
back to a CPU compute path.

  * `closed_loop`              main.py:121-152 tick loop (solve -> apply U[0:2] -> plant RK4 step) for a batch,
                               warm start unshifted like controller.py:57 (or shifted, §8(f)-1);
  * `event_triggered_resolve`  config 3: solve, perturb the state, re-solve warm ("re-solve sweep");
  * `predict_position`         event_triggered.py:18-28 `get_states` (a16), batched;
  * `obstacles_from_trajectories` / `best_response_rounds`
                               config 4: iterated best response of vehicle pairs; after every round each rank's
                               predicted trajectories [B_local, N, 4] are all-gathered (parallel.py) and become
                               the moving obstacles of the opponents' next solve.
"""
import torch

from . import parallel
from ._ffi import FLAG_BOUNDED_D


def predict_position(x, u_x, a, t0, t1):
    """event_triggered.py:26: SE(t0+t1) = x + u_x (t0+t1) + a (t0+t1)^2 / 2 (tensors or scalars)."""
    t = t0 + t1
    return x + u_x * t + 0.5 * a * t * t


def shift_warm_start(U, N):
    """Horizon shift of a flat `[d0, δ0, d1, δ1, ...]` warm start: drop stage 0, repeat the last stage."""
    B = U.shape[0]
    V = U.view(B, N, 2)
    return torch.cat([V[:, 1:], V[:, -1:]], dim=1).reshape(B, 2 * N).contiguous()


def closed_loop(prob, x0, road, ticks, road_idx=None, obs=None, U0=None, shift=False, sp=None, record_U=False):
    """main.py:121-152 for a batch, entirely on device.

    Per tick: `U = solve(y_n)` from the previous (U, λ) (controller.py:57, no shift unless `shift`), apply
    `U[:, 0:2]` to the plant for one RK4 stage of `Ts` (main.py:136-141: the plant is the model itself).
    Returns dict(states [ticks+1, B, nx], inputs [ticks, B, 2], status [ticks, B], inner [ticks, B],
    failures [B] (controller.py:64-66), U (last warm start))."""
    x = prob._t(x0).reshape(-1, prob.nx).clone()
    B, N = x.shape[0], prob.N
    dev = prob.device
    U = (torch.tensor([1.0, 0.0], dtype=torch.float64, device=dev).repeat(B, N) if U0 is None
         else prob._t(U0).reshape(B, 2 * N).clone())  # controller.py:20
    lam = torch.zeros(B, prob.m, dtype=torch.float64, device=dev) if prob.desc.flags & FLAG_BOUNDED_D else None
    states = torch.empty(ticks + 1, B, prob.nx, dtype=torch.float64, device=dev)
    inputs = torch.empty(ticks, B, 2, dtype=torch.float64, device=dev)
    status = torch.empty(ticks, B, dtype=torch.int32, device=dev)
    inner = torch.empty(ticks, B, dtype=torch.int32, device=dev)
    Us = torch.empty(ticks, B, 2 * N, dtype=torch.float64, device=dev) if record_U else None
    states[0] = x
    for t in range(ticks):
        U, lam, st = prob.solve(x, road, U, lam=lam, road_idx=road_idx, obs=obs, sp=sp)
        u = U[:, 0:2].contiguous()
        x = prob.plant_step(x, u)
        states[t + 1], inputs[t], status[t], inner[t] = x, u, st.status, st.inner_iterations
        if record_U:
            Us[t] = U
        if shift:
            U = shift_warm_start(U, N)
    out = dict(states=states, inputs=inputs, status=status, inner=inner, failures=(status != 1).sum(dim=0), U=U)
    if record_U:
        out["U_ticks"] = Us
    return out


def event_triggered_resolve(prob, x0, road, U0, road_idx=None, obs=None, sigma=0.02, seed=0, sp=None):
    """Config 3: the event trigger fires when the measured state leaves the predicted one, and the problem is
    re-solved from the previous solution.  Solves every instance cold, perturbs the state by N(0, sigma^2)
    (torch CPU generator -> reproducible on any device), re-solves warm.  Returns (U_cold, stats_cold,
    x_perturbed, U_warm, stats_warm)."""
    x0 = prob._t(x0).reshape(-1, prob.nx)
    U = prob._t(U0).reshape(x0.shape[0], prob.n).clone()
    lam = torch.zeros(x0.shape[0], prob.m, dtype=torch.float64, device=prob.device) if prob.desc.flags & FLAG_BOUNDED_D else None
    U, lam, st_cold = prob.solve(x0, road, U, lam=lam, road_idx=road_idx, obs=obs, sp=sp)
    g = torch.Generator().manual_seed(int(seed))
    noise = torch.randn(x0.shape, dtype=torch.float64, generator=g).to(prob.device) * sigma
    x1 = (x0 + noise).contiguous()
    U_cold = U.clone()
    Uw, lam, st_warm = prob.solve(x1, road, U, lam=lam, road_idx=road_idx, obs=obs, sp=sp)
    return U_cold, st_cold, x1, Uw, st_warm


def obstacles_from_trajectories(x_now, traj, Ts):
    """Obstacle records [B, 1, 5] = (x, y, φ, v, a) for the kernel's constant-acceleration obstacle motion
    (event_triggered.py:26) from the opponents' current states `x_now` [B, 4] and predicted trajectories
    `traj` [B, N, 4] (x, y, φ, v per stage): a = mean acceleration over the horizon."""
    N = traj.shape[1]
    a = (traj[:, -1, 3] - x_now[:, 3]) / (N * Ts)
    return torch.stack([x_now[:, 0], x_now[:, 1], x_now[:, 2], x_now[:, 3], a], dim=1).unsqueeze(1).contiguous()


def opponent_index(n_vehicles, device=None):
    """Vehicles are stored [A_0..A_{P-1}, B_0..B_{P-1}]; the opponent of vehicle v is (v + P) mod 2P."""
    P = n_vehicles // 2
    return (torch.arange(n_vehicles, device=device) + P) % n_vehicles


def lane_of(y, h=3.75):
    """Lane index (1 or 2, as in game_theory.py's Car.lane) of a lateral position: lane 1 is centred at y = 0, lane 2 at h."""
    return torch.where(y < 0.5 * h, torch.ones_like(y), torch.full_like(y, 2.0))


def decide_paths(prob, x_now, lead, opp_now, opp_lane, N, h=3.75):
    """The game-theory layer of config 4 on the device (mpcb_game_payoff = game_theory.py:205-244 + :92-113): every vehicle
    scores target lanes 1 and 2 against its scenario's lead car and its opponent (current position and speed, in the lane
    the opponent ANNOUNCED — the lane its last gathered trajectory ends in), takes the better one and gets the Bezier
    control points of its path: `get_bezier_control_points` towards the car in front for a lane change (mirrored for
    2 -> 1), a straight path in its own lane otherwise (or when nobody is in front).
    x_now [B,4] = (x, y, phi, v); lead [B,3]; opp_now [B,4]; opp_lane [B].  Returns (ctrl [B,12] world frame, target [B],
    payoff [B,2])."""
    B = x_now.shape[0]
    lane = lane_of(x_now[:, 1], h)
    ego = torch.stack([x_now[:, 0], x_now[:, 3], lane], dim=1).contiguous()
    cars = torch.stack([lead, torch.stack([opp_now[:, 0], opp_now[:, 3], opp_lane], dim=1)], dim=1).contiguous()
    payoff, decision, rel, _ = prob.game_payoff(ego, cars)
    target = torch.where(decision > 0, torch.full_like(lane, 2.0), torch.ones_like(lane))
    change = (target != lane) & torch.isfinite(rel).all(dim=1) & (rel[:, 5] > 1.0)
    # straight path in the own lane, long enough for the horizon
    Lh = torch.clamp(x_now[:, 3], min=1.0) * (N * prob.desc.Ts) * 2.5
    frac = torch.linspace(0.0, 1.0, 6, dtype=torch.float64, device=x_now.device)
    y_own = (lane - 1.0) * h
    straight = torch.cat([x_now[:, 0:1] + Lh[:, None] * frac[None], y_own[:, None].expand(B, 6)], dim=1)
    py = torch.where((lane == 1.0)[:, None], rel[:, 6:], h - rel[:, 6:])   # 1 -> 2 as the reference, 2 -> 1 mirrored
    moved = torch.cat([x_now[:, 0:1] + rel[:, :6], py], dim=1)
    ctrl = torch.where(change[:, None], moved, straight).contiguous()
    return ctrl, target, payoff


def best_response_rounds(prob, x0_local, ctrl_local, n_vehicles, rounds=5, U0=None, group=None, rank=None,
                         world=None, sp=None, cars=None, timing=False):
    """Config 4: iterated best response of 2-vehicle scenarios (game_theory.py roles: each vehicle plans its lane
    change against the other's announced plan).

    `x0_local` [B_local, 4]: this rank's contiguous shard (parallel.shard_range) of the 2P vehicles.  Every round: (with
    `cars`, the dict of workloads.game_pairs) the payoff layer picks every local vehicle's lane and Bezier path
    (`decide_paths`); solve the local vehicles against the current obstacle records; roll the solutions out; ALL-GATHER
    the predicted trajectories `[B_local, N, 4]` over the group (NCCL on GPUs) and rebuild each vehicle's obstacle from
    its opponent's row.  Without `cars` the paths are the fixed `ctrl_local` [B_local, 12] (no decision layer).
    Returns dict(U, traj (global [2P, N, 4]), status [rounds, B_local], delta [rounds] = max change of the gathered
    trajectories between rounds (delta_median: the median over vehicles of each vehicle's max change), obs; with `cars`: target [rounds, B_local], payoff, lane_change_frac; with `timing`:
    solve_ms / gather_ms per round from CUDA events)."""
    import torch.distributed as dist
    if rank is None:
        rank = dist.get_rank(group) if dist.is_available() and dist.is_initialized() else 0
    if world is None:
        world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    dev = prob.device
    x0 = prob._t(x0_local).reshape(-1, prob.nx)
    Bl, N = x0.shape[0], prob.N
    lo, hi = parallel.shard_range(n_vehicles, rank, world)
    assert hi - lo == Bl, "x0_local must be this rank's shard of the vehicle list"
    opp = opponent_index(n_vehicles, dev)[lo:hi]
    U = torch.zeros(Bl, 2 * N, dtype=torch.float64, device=dev) if U0 is None else prob._t(U0).clone()
    x_all = parallel.allgather_trajectories(x0, n_vehicles, group=group)
    # round 0 announcement: every opponent keeps its speed (a = 0) and its lane
    obs = obstacles_from_trajectories(x_all[opp], x_all[opp].unsqueeze(1).expand(-1, N, -1), prob.desc.Ts)
    game = cars is not None
    if game:
        h = float(cars.get("h", 3.75))
        lead = prob._t(cars["lead"])[lo:hi].contiguous()
        opp_lane = lane_of(x_all[opp][:, 1], h)
        # the lead car is the second obstacle of every solve: (x, y, phi, v, a) on the lane-1 centre line, constant speed
        lead_obs = torch.stack([lead[:, 0], torch.zeros_like(lead[:, 0]), torch.zeros_like(lead[:, 0]), lead[:, 1],
                                torch.zeros_like(lead[:, 0])], dim=1).unsqueeze(1)
        ctrl = None
    else:
        ctrl = prob._t(ctrl_local).reshape(-1, 12)
    status, delta, delta_med, prev, targets, payoff = [], [], [], None, [], None
    traj_all = None
    ev = []
    for _ in range(rounds):
        if game:
            ctrl, target, payoff = decide_paths(prob, x0, lead, x_all[opp], opp_lane, N, h)
            targets.append(target)
            obs_in = torch.cat([obs, lead_obs], dim=1).contiguous()
        else:
            obs_in = obs
        if timing:
            e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            e[0].record()
        U, _, st = prob.solve(x0, ctrl, U, obs=obs_in, sp=sp)
        traj, _, _ = prob.rollout(x0, U, ctrl, obs=obs_in)
        if timing:
            e[1].record()
        traj_all = parallel.allgather_trajectories(traj.contiguous(), n_vehicles, group=group)
        if timing:
            e[2].record()
            ev.append(e)
        obs = obstacles_from_trajectories(x_all[opp], traj_all[opp], prob.desc.Ts)
        if game:
            opp_lane = lane_of(traj_all[opp][:, -1, 1], h)
        status.append(st.status)
        delta.append(float("inf") if prev is None else float((traj_all - prev).abs().max().item()))
        delta_med.append(float("inf") if prev is None else
                         float((traj_all - prev).abs().flatten(1).max(dim=1).values.median().item()))
        prev = traj_all
    out = dict(U=U, traj=traj_all, status=torch.stack(status), delta=delta, delta_median=delta_med, obs=obs)
    if game:
        tg = torch.stack(targets)
        out.update(target=tg, payoff=payoff, ctrl=ctrl,
                   lane_change_frac=float((tg[-1] != lane_of(x0[:, 1], h)).double().mean().item()))
    if timing:
        torch.cuda.synchronize()
        out["solve_ms"] = [e[0].elapsed_time(e[1]) for e in ev]
        out["gather_ms"] = [e[1].elapsed_time(e[2]) for e in ev]
    return out
