"""Evaluation rounds per instance (MPCB_DEBUG_TIMERS=3 puts them in the psi output): where the critical path goes."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["MPCB_DEBUG_TIMERS"] = "3"


def main():
    import numpy as np
    import torch
    import mpcb200
    from mpcb200 import workloads as wl
    B = int(os.environ.get("PROF_B", "4096"))
    N = int(os.environ.get("PROF_N", "20"))
    w = wl.tracking_batch(B, N, 100, seed=2)
    prob = mpcb200.BatchedMpc(N=N, S=100, road_mode=mpcb200.ROAD_TABLE, n_roads=2)
    dev = prob.device
    x0, road = torch.tensor(w["x0"], device=dev), torch.tensor(w["road"], device=dev)
    ridx, U = torch.tensor(w["road_idx"], device=dev), torch.tensor(w["U0"], device=dev)
    _, _, st = prob.solve(x0, road, U, road_idx=ridx)
    torch.cuda.synchronize()
    packed = st.psi.cpu().numpy()
    slides = np.floor(packed / 1e12); packed -= slides * 1e12
    lemma = np.floor(packed / 1e8); packed -= lemma * 1e8
    fdr = np.floor(packed / 1e4); rounds = packed - fdr * 1e4
    it = st.inner_iterations.cpu().numpy().astype(float)
    outer = st.outer_iterations.cpu().numpy().astype(float)
    print(f"B={B}: rounds mean {rounds.mean():.1f} max {rounds.max():.0f}; iterations mean {it.mean():.1f}; outer mean {outer.mean():.2f}; "
          f"rounds/iteration overall {rounds.sum() / it.sum():.3f}; (rounds - 6*outer)/iteration {((rounds - 6 * outer).sum()) / it.sum():.3f}")
    o = np.argsort(-rounds)[:12]
    for i in o:
        print(f"  instance {i}: rounds {rounds[i]:.0f} iterations {it[i]:.0f} outer {outer[i]:.0f} ratio {rounds[i] / it[i]:.2f} "
              f"evals v {st.evals_value[i].item()} g {st.evals_grad[i].item()}; FD rounds {fdr[i]:.0f}, lemma re-evaluations (any lane) {lemma[i]:.0f}, slides {slides[i]:.0f}")
    print(f"totals per instance: FD rounds {fdr.mean():.1f}, lemma re-evaluations {lemma.mean():.1f}, slides {slides.mean():.1f}")


if __name__ == "__main__":
    main()
