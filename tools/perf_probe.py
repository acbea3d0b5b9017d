"""Performance probe (run on the GPU box): dependent-chain latencies, and rollout / value_grad / solve
throughput as a function of batch size.  Writes gpurun_out/perf_probe.json."""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timed(fn, reps=3):
    import torch
    fn()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best


def main():
    import torch
    import mpcb200
    from mpcb200 import workloads as wl
    out = {}
    lat = (C.c_double * 8)()
    mpcb200._ffi.check(mpcb200._ffi.lib().mpcb_latency_probe(lat), None, "latency_probe")
    out["latency_cycles"] = dict(zip(["dfma", "div", "sqrt", "sincos", "atan", "atan2", "sin", "pacejka_f"], list(lat)))
    out["fp64_peak_tflops"] = mpcb200.fp64_peak_tflops()[0]
    print(json.dumps(out), flush=True)
    N = 20
    sizes = [int(s) for s in os.environ.get("PROBE_SIZES", "32,1024,4096,16384,65536,262144").split(",")]
    prob = mpcb200.BatchedMpc(N=N, S=100, road_mode=mpcb200.ROAD_TABLE, n_roads=2)
    dev = prob.device
    rows = []
    for B in sizes:
        w = wl.tracking_batch(B, N, 100, seed=2)
        x0, road = torch.tensor(w["x0"], device=dev), torch.tensor(w["road"], device=dev)
        ridx, U0 = torch.tensor(w["road_idx"], device=dev), torch.tensor(w["U0"], device=dev)
        t_roll = timed(lambda: prob.rollout(x0, U0, road, road_idx=ridx))
        t_vg = timed(lambda: prob.value_grad(x0, U0, road, road_idx=ridx))
        row = {"B": B, "rollout_ms": t_roll, "value_grad_ms": t_vg, "rollout_per_s": B / t_roll * 1e3,
               "value_grad_per_s": B / t_vg * 1e3}
        if B <= int(os.environ.get("PROBE_SOLVE_MAX", "65536")):
            Us = [U0.clone() for _ in range(3)]
            it = iter(Us)
            t_solve = timed(lambda: prob.solve(x0, road, next(it), road_idx=ridx), reps=2)
            row.update(solve_ms=t_solve, solves_per_s=B / t_solve * 1e3)
        rows.append(row)
        print(json.dumps(row), flush=True)
    out["scaling"] = rows
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "perf_probe.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
