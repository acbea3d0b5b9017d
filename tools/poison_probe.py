"""Do results depend on memory a solve did not write?  Same batch with MPCB_POISON=0/1 and several speculation widths:
every pair must be bit-identical (diagnostic for uninitialised reads; GPU box)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import mpcb200 as mp
    from mpcb200 import workloads as wl
    os.environ["MPCB_MODE"] = "rounds"
    B, N = int(os.environ.get("PB", "4096")), 20
    w = wl.tracking_batch(B, N, 100, seed=2)
    res = {}
    for poison in ("0", "1"):
        for W in ("7", "3", "1"):
            os.environ["MPCB_POISON"], os.environ["MPCB_SPECW"] = poison, W
            prob = mp.BatchedMpc(N=N, S=100, road_mode=mp.ROAD_TABLE, n_roads=2)
            U, _, st = prob.solve(w["x0"], w["road"], w["U0"].copy(), road_idx=w["road_idx"])
            torch.cuda.synchronize()
            res[(poison, W)] = (U.clone(), st.inner_iterations.clone())
            del prob
    ref = res[("0", "7")]
    for k, (U, it) in res.items():
        same = (U == ref[0]).all(dim=1)
        nan = torch.isnan(U).any(dim=1)
        print(f"poison={k[0]} W={k[1]}: bit-identical to (0,7) {same.double().mean().item():.4f}, NaN rows {int(nan.sum())}, "
              f"inner counts equal {(it == ref[1]).double().mean().item():.4f}", flush=True)


if __name__ == "__main__":
    main()
