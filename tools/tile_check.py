"""Solve the same batch with every tile width (MPCB_TILE = 8, 4, 2, 1) and compare iterations and solutions:
the widths differ only in how much of the line search / FD gradient is speculated, not in the decisions."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import numpy as np
    import torch
    import mpcb200
    from mpcb200 import workloads as wl
    B = int(os.environ.get("PROF_B", "2048"))
    N = int(os.environ.get("PROF_N", "20"))
    w = wl.tracking_batch(B, N, 100, seed=2)
    ref = None
    for G in (8, 4, 2, 1):
        os.environ["MPCB_TILE"] = str(G)
        prob = mpcb200.BatchedMpc(N=N, S=100, road_mode=mpcb200.ROAD_TABLE, n_roads=2)
        dev = prob.device
        x0, road = torch.tensor(w["x0"], device=dev), torch.tensor(w["road"], device=dev)
        ridx = torch.tensor(w["road_idx"], device=dev)
        for rep in range(2):
            U = torch.tensor(w["U0"], device=dev)
            torch.cuda.synchronize()
            t = time.perf_counter()
            _, _, st = prob.solve(x0, road, U, road_idx=ridx)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t
        it = st.inner_iterations.cpu().numpy()
        Uh = U.cpu().numpy()
        line = (f"G={G}: {dt * 1e3:.1f} ms, converged {(st.status == 1).float().mean().item():.4f}, inner mean {it.mean():.2f} "
                f"max {it.max()}, evals value {st.evals_value.double().mean().item():.1f} grad {st.evals_grad.double().mean().item():.1f}")
        if ref is None:
            ref = (it, Uh)
        else:
            line += (f"; vs G=8: iterations differ in {(it != ref[0]).mean():.4f} of instances, "
                     f"max |dU| {np.abs(Uh - ref[1]).max():.2e}")
        print(line, flush=True)


if __name__ == "__main__":
    main()
