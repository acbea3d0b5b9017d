"""Is the evaluation time data-dependent?  value_grad / rollout over config-2 instances with the cold-start inputs
(U0 = [1, 0] tiled) against random inputs in the box and against mid-solve iterates (the solution itself)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import mpcb200
    from mpcb200 import workloads as wl
    N = 20
    prob = mpcb200.BatchedMpc(N=N, S=100, road_mode=mpcb200.ROAD_TABLE, n_roads=2)
    dev = prob.device
    for B in (4096, 16384):
        w = wl.tracking_batch(B, N, 100, seed=2)
        wr = wl.tracking_batch(B, N, 100, seed=2, random_U=True)
        x0, road = torch.tensor(w["x0"], device=dev), torch.tensor(w["road"], device=dev)
        ridx = torch.tensor(w["road_idx"], device=dev)
        U0, Ur = torch.tensor(w["U0"], device=dev), torch.tensor(wr["U0"], device=dev)
        Us, _, _ = prob.solve(x0, road, U0.clone(), road_idx=ridx)
        for name, U in (("cold U0", U0), ("random U", Ur), ("solution U*", Us)):
            for what, fn in (("value_grad", lambda: prob.value_grad(x0, U, road, road_idx=ridx)),
                             ("rollout", lambda: prob.rollout(x0, U, road, road_idx=ridx))):
                fn(); torch.cuda.synchronize()
                best = 1e9
                for _ in range(3):
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record(); fn(); b.record(); torch.cuda.synchronize()
                    best = min(best, a.elapsed_time(b))
                print(f"B={B} {name:12s} {what:10s} {best:.3f} ms", flush=True)


if __name__ == "__main__":
    main()
