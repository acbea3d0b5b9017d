"""Solve throughput by execution mode, tile width and batch size (run on the GPU box).
PROBE="mode:G:B[:W],..." e.g. "rounds:8:4096,persistent:8:4096,rounds:8:65536:1" (W = MPCB_SPECW).  One subprocess-free pass: the env overrides are
read at mpcb_create, so every combination gets its own handle."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import mpcb200
    from mpcb200 import workloads as wl
    N = int(os.environ.get("PROBE_N", "20"))
    combos = os.environ.get("PROBE", "rounds:8:4096,persistent:8:4096").split(",")
    rows = []
    for c in combos:
        f = c.split(":")
        mode, G, B = f[0], f[1], int(f[2])
        W = f[3] if len(f) > 3 else "0"   # speculation window: 0 = device policy, 1..7 fixed (MPCB_SPECW)
        if W != "0":
            os.environ["MPCB_SPECW"] = W
        else:
            os.environ.pop("MPCB_SPECW", None)
        os.environ["MPCB_MODE"] = mode
        if int(G) > 0:
            os.environ["MPCB_TILE"] = G
        else:
            os.environ.pop("MPCB_TILE", None)
        prob = mpcb200.BatchedMpc(N=N, S=100, road_mode=mpcb200.ROAD_TABLE, n_roads=2)
        dev = prob.device
        w = wl.tracking_batch(B, N, 100, seed=int(os.environ.get("PROBE_SEED", "2")))
        x0, road = torch.tensor(w["x0"], device=dev), torch.tensor(w["road"], device=dev)
        ridx, U0 = torch.tensor(w["road_idx"], device=dev), torch.tensor(w["U0"], device=dev)
        best, st = 1e30, None
        for i in range(int(os.environ.get("PROBE_REPS", "2")) + 1):
            U = U0.clone()
            l0 = prob.launch_count()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            _, _, st = prob.solve(x0, road, U, road_idx=ridx)
            b.record()
            torch.cuda.synchronize()
            if i > 0:
                best = min(best, a.elapsed_time(b))
            launches = prob.launch_count() - l0
        row = dict(mode=mode, G=int(G), B=B, W=int(W), ms=best, solves_per_s=B / best * 1e3, launches=launches,
                   converged=float((st.status == 1).double().mean().item()),
                   inner_mean=float(st.inner_iterations.double().mean().item()),
                   inner_max=int(st.inner_iterations.max().item()),
                   evals_value=float(st.evals_value.double().mean().item()),
                   evals_grad=float(st.evals_grad.double().mean().item()), profile=prob.last_solve_profile())
        rows.append(row)
        print(json.dumps(row), flush=True)
        del prob
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(rows, open(os.path.join(ROOT, "gpurun_out", os.environ.get("PROBE_OUT", "mode_probe.json")), "w"), indent=1)


if __name__ == "__main__":
    main()
