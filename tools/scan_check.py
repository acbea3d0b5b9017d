"""Pruned nearest-vertex search (rollout kernel, hinted, shared-memory road) against the full scan
(road_errors kernel) on the rollout's own stage positions — every index must agree — and the heading /
wrap path through the rollout cost against the CPU oracle on a sample."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import numpy as np
    import torch
    import mpcb200
    from mpcb200 import workloads as wl
    from oracle import oracle as orc
    B, N, S = int(os.environ.get("PROF_B", "16384")), 20, 100
    w = wl.tracking_batch(B, N, S, seed=3)
    rng = np.random.default_rng(0)
    U = w["U0"] + rng.uniform(-0.3, 0.3, w["U0"].shape) * np.tile([1.0, 0.3], N)
    prob = mpcb200.BatchedMpc(N=N, S=S, road_mode=mpcb200.ROAD_TABLE, n_roads=2)
    states, idx, cost = prob.rollout(w["x0"], U, w["road"], road_idx=w["road_idx"])
    pos = states[:, :, :2].reshape(-1, 2).contiguous()
    hd = states[:, :, 2].reshape(-1).contiguous()
    ridx = torch.tensor(np.repeat(w["road_idx"], N).astype(np.int32), device=prob.device)
    idx2, err = prob.road_errors(pos, hd, w["road"], road_idx=ridx)
    bad = (idx.reshape(-1) != idx2).sum().item()
    print(f"indices compared {idx.numel()}, mismatches {bad}")
    d = orc.default_desc(0, N, S); d.road_mode, d.n_roads = 1, 2
    n = 2048
    r = orc.rollout(d, w["x0"][:n], U[:n], w["road"], road_idx=w["road_idx"][:n])
    c_ref = r[2] if isinstance(r, tuple) else r["cost"]
    rel = np.abs(cost[:n].cpu().numpy() - c_ref) / np.abs(c_ref)
    print(f"rollout cost vs oracle: max rel {rel.max():.2e}")


if __name__ == "__main__":
    main()
