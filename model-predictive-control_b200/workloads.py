"""Synthetic workloads of SURVEY.md §8(d) (numpy only; RNG = PCG64(seed)).

Roads follow reference main.py:11-22 (`get_centerline`); initial states follow the config-2 recipe:
start vertex k ~ U{5..90}, lateral offset, heading offset, vx in [0.5, 1.5] (the Pacejka model is stiff
below vx ~ 0.3), small vy / omega.
"""
import numpy as np


def get_centerline(size, is_straight=True):
    """reference main.py:11-22, returned as the (size, 2) array the reference builds."""
    if is_straight:
        return np.array([[i / 10 - 0.1, 0] for i in range(size)], dtype=np.float64)
    theta = np.linspace(0, 2 * np.pi, size)
    x = 5 * np.cos(theta)
    y = 5 * np.sin(theta) + 5
    return np.stack((x, y), axis=1)


def flat_F(centerline):
    """(S,2) -> flat column-major [x0..xS-1, y0..yS-1] (main.py:113)."""
    return np.ascontiguousarray(centerline.ravel(order="F"))


def road_table(S=100):
    """[2, 2S]: row 0 straight, row 1 circle."""
    return np.stack([flat_F(get_centerline(S, True)), flat_F(get_centerline(S, False))])


def tracking_batch(B, N=20, S=100, seed=2, lateral=0.05, model="pacejka", random_U=False, d_min=-1.0):
    """Config 2: B independent tracking instances on a 2-road table.

    Returns dict(x0 [B,nx], road [2,2S], road_idx [B] int32, U0 [B,2N])."""
    rng = np.random.Generator(np.random.PCG64(seed))
    table = road_table(S)
    ridx = rng.integers(0, 2, B).astype(np.int32)
    k = rng.integers(5, 91, B)
    e = rng.uniform(-lateral, lateral, B)
    dphi = rng.uniform(-0.3, 0.3, B)
    vx = rng.uniform(0.5, 1.5, B)
    vy = rng.uniform(-0.05, 0.05, B)
    om = rng.uniform(-0.5, 0.5, B)
    xs = table[ridx, :S]
    ys = table[ridx, S:]
    ar = np.arange(B)
    px, py = xs[ar, k], ys[ar, k]
    tx, ty = xs[ar, k + 1] - px, ys[ar, k + 1] - py
    tn = np.hypot(tx, ty)
    tx, ty = tx / tn, ty / tn
    pos_x = px - ty * e
    pos_y = py + tx * e
    phi = np.arctan2(ty, tx) + dphi
    if model == "pacejka":
        x0 = np.stack([pos_x, pos_y, phi, vx, vy, om], axis=1)
    else:
        x0 = np.stack([pos_x, pos_y, phi, vx], axis=1)
    if random_U:
        U0 = np.empty((B, 2 * N))
        U0[:, 0::2] = rng.uniform(d_min, 1, (B, N))  # d_min > 0 keeps vx out of the stiff low-speed regime
        U0[:, 1::2] = rng.uniform(-0.32, 0.32, (B, N))
    else:
        U0 = np.tile([1.0, 0.0], (B, N))  # controller.py:20
    return dict(x0=np.ascontiguousarray(x0), road=table, road_idx=ridx, U0=np.ascontiguousarray(U0))


def lane_change_batch(B, N=40, S=100, seed=3, n_obs=2):
    """Config 3/4 style full-scale instances (Kinematic-4): Bezier lane-change roads
    (bezier_curves.py:28-48 geometry with randomised gap/speeds) and n_obs moving obstacles.

    Returns dict(x0 [B,4], ctrl [B,12], obs [B,n_obs,5], U0 [B,2N])."""
    rng = np.random.Generator(np.random.PCG64(seed))
    h, W, theta, l, Lf = 3.75, 1.8, 3.2 / 180 * np.pi, 3.0, 1.0
    v0 = rng.uniform(15, 25, B)
    v1 = v0 - rng.uniform(5, 12, B)
    D1 = rng.uniform(35, 70, B)
    i = 5.0
    Di = (Lf + l) * np.cos(np.arctan2(W, 2 * Lf) - theta)
    tc1 = D1 / (v0 - v1)
    Px2 = v0 * tc1 - Di
    Px5 = 2 * Px2
    ctrl = np.zeros((B, 12))
    ctrl[:, 1] = Px2 / i
    ctrl[:, 2] = Px2
    ctrl[:, 3] = Px2
    ctrl[:, 4] = Px5 - (Px5 - Px2) / i
    ctrl[:, 5] = Px5
    ctrl[:, 9:12] = h
    x0 = np.stack([rng.uniform(0, 2, B), rng.uniform(-0.2, 0.2, B), rng.uniform(-0.02, 0.02, B), v0], axis=1)
    obs = np.zeros((B, n_obs, 5))
    for o in range(n_obs):
        obs[:, o, 0] = D1 * (1 + o) + rng.uniform(-2, 2, B)   # x
        obs[:, o, 1] = 0.0 if o == 0 else h                   # lane
        obs[:, o, 2] = 0.0
        obs[:, o, 3] = v1 * rng.uniform(0.9, 1.1, B)
        obs[:, o, 4] = rng.uniform(-0.5, 0.5, B)
    U0 = np.tile([0.0, 0.0], (B, N))
    return dict(x0=np.ascontiguousarray(x0), ctrl=ctrl, obs=obs, U0=np.ascontiguousarray(U0))


def game_pairs(P, seed=4, h=3.75):
    """Config 4: P two-vehicle lane-change scenarios around reference game_theory.py:247-349 (`get_cars_test_1..3`: ego in
    lane 1 closing in on a slow car, faster traffic in lane 2), scaled to the Kinematic-4 lane-change setting.

    Vehicles are stored [A_0..A_{P-1}, B_0..B_{P-1}] (opponent of v is (v + P) mod 2P):  A_i drives in lane 1 (y = 0) behind
    a slow lead car; B_i drives in lane 2 (y = h) behind A_i and faster.  Returns dict(x0 [2P,4] = (x, y, phi, v),
    lead [2P,3] = (x, v, lane) of each vehicle's scenario lead car, U0 [2P, 2N] is left to the caller)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    vA = rng.uniform(15, 25, P)
    xA = rng.uniform(0, 2, P)
    lead_x = xA + rng.uniform(35, 70, P)
    lead_v = vA - rng.uniform(5, 12, P)
    xB = xA - rng.uniform(8, 30, P)
    vB = vA + rng.uniform(-2, 5, P)
    x0 = np.zeros((2 * P, 4))
    x0[:P] = np.stack([xA, rng.uniform(-0.2, 0.2, P), rng.uniform(-0.02, 0.02, P), vA], axis=1)
    x0[P:] = np.stack([xB, h + rng.uniform(-0.2, 0.2, P), rng.uniform(-0.02, 0.02, P), vB], axis=1)
    lead = np.stack([lead_x, lead_v, np.ones(P)], axis=1)
    return dict(x0=np.ascontiguousarray(x0), lead=np.ascontiguousarray(np.concatenate([lead, lead])), h=h)
