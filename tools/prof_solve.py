"""Profiling target: config-2 solves (B instances, N=20) — one warm-up and one measured launch of k_solve."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import mpcb200
    from mpcb200 import workloads as wl
    B = int(os.environ.get("PROF_B", "4096"))
    N = int(os.environ.get("PROF_N", "20"))
    prob = mpcb200.BatchedMpc(N=N, S=100, road_mode=mpcb200.ROAD_TABLE, n_roads=2)
    dev = prob.device
    w = wl.tracking_batch(B, N, 100, seed=2)
    x0, road = torch.tensor(w["x0"], device=dev), torch.tensor(w["road"], device=dev)
    ridx, U0 = torch.tensor(w["road_idx"], device=dev), torch.tensor(w["U0"], device=dev)
    for i in range(int(os.environ.get("PROF_REPS", "2"))):
        U = U0.clone()
        torch.cuda.synchronize()
        t = time.perf_counter()
        _, _, st = prob.solve(x0, road, U, road_idx=ridx)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t
        print(f"solve {i}: {dt * 1e3:.1f} ms, {B / dt:.0f} solves/s, converged {(st.status == 1).float().mean().item():.4f}, "
              f"inner mean {st.inner_iterations.double().mean().item():.1f} max {st.inner_iterations.max().item()}, "
              f"evals value {st.evals_value.double().mean().item():.1f} grad {st.evals_grad.double().mean().item():.1f}")
    if os.environ.get("MPCB_DEBUG_TIMERS"):
        import ctypes as C
        import numpy as np
        buf = (C.c_longlong * (4 * 4096))()
        nw = mpcb200._ffi.lib().mpcb_debug_timers(prob._h, buf, 4096)
        if nw == 0:
            return
        t = np.frombuffer(buf, dtype=np.int64)[:4 * nw].reshape(nw, 4).astype(float)
        tot = t[:, 0] + t[:, 1]
        i = int(np.argmax(tot))
        print(f"timers over {nw} warps: eval share {t[:, 0].sum() / tot.sum():.3f}, mean rounds {t[:, 2].mean():.0f} (grad {t[:, 3].mean():.0f}), "
              f"cycles/round eval {t[:, 0].sum() / t[:, 2].sum():.0f} transition {t[:, 1].sum() / t[:, 2].sum():.0f}; "
              f"slowest warp: {tot[i] / 1.965e6:.1f} ms, rounds {t[i, 2]:.0f}, eval/round {t[i, 0] / t[i, 2]:.0f}, trans/round {t[i, 1] / t[i, 2]:.0f}")


if __name__ == "__main__":
    main()
