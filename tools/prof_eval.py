"""Profiling target: one rollout and one value_grad launch over config-2 style instances (run under ncu)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import mpcb200
    from mpcb200 import workloads as wl
    B = int(os.environ.get("PROF_B", "65536"))
    N = 20
    prob = mpcb200.BatchedMpc(N=N, S=100, road_mode=mpcb200.ROAD_TABLE, n_roads=2)
    dev = prob.device
    w = wl.tracking_batch(B, N, 100, seed=2)
    x0, road = torch.tensor(w["x0"], device=dev), torch.tensor(w["road"], device=dev)
    ridx, U0 = torch.tensor(w["road_idx"], device=dev), torch.tensor(w["U0"], device=dev)
    for _ in range(2):
        prob.rollout(x0, U0, road, road_idx=ridx)
        prob.value_grad(x0, U0, road, road_idx=ridx)
    torch.cuda.synchronize()
    print("ok")


if __name__ == "__main__":
    main()
