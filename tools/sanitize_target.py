"""Small target for compute-sanitizer (memcheck / racecheck / synccheck, ONE tool per gpurun call): every kernel family on
small batches — both execution modes, tile widths 8 and 1, narrow and wide speculation windows, split evaluation on,
bounded-D ALM, Kinematic-4 with obstacles on Bezier roads, value_grad (allocated and static scratch), the payoff kernel."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import numpy as np
    import torch
    import mpcb200 as mp
    from mpcb200 import workloads as wl
    N, B = 12, int(os.environ.get("SAN_B", "48"))
    w = wl.tracking_batch(B, N, 100, seed=3)
    ref = None
    for mode in ("rounds", "persistent"):
        for tile, specw in (("8", "7"), ("8", "1"), ("1", "1")):
            os.environ.update(MPCB_MODE=mode, MPCB_TILE=tile, MPCB_SPECW=specw, MPCB_SPLIT="1")
            prob = mp.BatchedMpc(N=N, S=100, road_mode=mp.ROAD_TABLE, n_roads=2)
            U, _, st = prob.solve(w["x0"], w["road"], w["U0"].copy(), road_idx=w["road_idx"])
            torch.cuda.synchronize()
            ok = float((st.status == 1).double().mean().item())
            if mode == "rounds":
                ref = U.clone() if ref is None else ref
                assert torch.equal(U, ref), "tile / window width changed the iterates"
            print(f"{mode} G={tile} W={specw}: converged {ok:.3f}, inner mean {st.inner_iterations.double().mean().item():.1f}", flush=True)
            del prob
    for k in ("MPCB_TILE", "MPCB_SPECW"):
        os.environ.pop(k, None)
    os.environ["MPCB_MODE"] = "rounds"
    # bounded-D ALM (multipliers, per-row penalties)
    prob = mp.BatchedMpc(N=N, S=100, road_mode=mp.ROAD_TABLE, n_roads=2, flags=mp.FLAG_BOUNDED_D, D_lower=[-np.inf] * 6,
                         D_upper=[np.inf, np.inf, np.inf, -1.0, np.inf, np.inf])
    U, lam, st = prob.solve(w["x0"], w["road"], w["U0"].copy(), lam=np.zeros((B, 6 * N)), road_idx=w["road_idx"])
    torch.cuda.synchronize()
    print(f"bounded D: statuses {np.bincount(st.status.cpu().numpy())}", flush=True)
    # Kinematic-4, Bezier roads, lane potentials, moving obstacles
    lc = wl.lane_change_batch(B, N, 100, seed=5, n_obs=2)
    probk = mp.BatchedMpc(N=N, S=100, model_id=1, road_mode=mp.ROAD_BEZIER, flags=mp.FLAG_LANE | mp.FLAG_OBST, n_obs=2, v_ref=20.0)
    U, _, st = probk.solve(lc["x0"], lc["ctrl"], lc["U0"].copy(), obs=lc["obs"])
    torch.cuda.synchronize()
    print(f"kinematic + obstacles: converged {float((st.status == 1).double().mean().item()):.3f}", flush=True)
    # evaluation entry points
    prob = mp.BatchedMpc(N=N, S=100, road_mode=mp.ROAD_TABLE, n_roads=2)
    dev = prob.device
    t = {k: torch.tensor(w[k], device=dev) for k in ("x0", "U0", "road", "road_idx")}
    c, g = prob.value_grad(t["x0"], t["U0"], t["road"], road_idx=t["road_idx"])
    c2, g2 = torch.empty_like(c), torch.empty_like(g)
    prob.value_grad_static(t["x0"], t["U0"], t["road"], c2, g2, road_idx=t["road_idx"])
    prob.rollout(t["x0"], t["U0"], t["road"], road_idx=t["road_idx"])
    rng = np.random.default_rng(1)
    ego = np.stack([rng.uniform(-5, 5, B), rng.uniform(3, 25, B), rng.integers(1, 3, B)], axis=1).astype(float)
    cars = np.stack([rng.uniform(-80, 100, (B, 4)), rng.uniform(0, 28, (B, 4)), rng.integers(0, 3, (B, 4))], axis=2).astype(float)
    prob.game_payoff(ego, cars)
    torch.cuda.synchronize()
    assert torch.equal(c, c2) and torch.equal(g, g2)
    print("sanitize target done", flush=True)


if __name__ == "__main__":
    main()
