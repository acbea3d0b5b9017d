"""BASELINE configs 3, 4, 5 on one GPU (per-GPU shares of the 8-GPU configurations); writes gpurun_out/configs.json.
Not the headline bench (bench.py): context numbers for DESIGN.md / profiles/."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
FLOP_PER_STAGE_GRAD = 3864.0


def ev_time(fn, reps=3):
    import torch
    fn(); torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best


def main():
    import torch
    import mpcb200
    from mpcb200 import scenarios as sc, workloads as wl
    out = {}
    peak, _ = mpcb200.fp64_peak_tflops()
    out["fp64_peak_tflops"] = peak
    # ---- config 5: 1M-instance rollout + gradient sweep (one GPU takes the whole sweep) ----
    B, N = 1 << 20, 20
    prob = mpcb200.BatchedMpc(N=N, S=100, road_mode=mpcb200.ROAD_TABLE, n_roads=2)
    dev = prob.device
    w = wl.tracking_batch(B, N, 100, seed=5, random_U=True)
    x0, road = torch.tensor(w["x0"], device=dev), torch.tensor(w["road"], device=dev)
    ridx, U = torch.tensor(w["road_idx"], device=dev), torch.tensor(w["U0"], device=dev)
    t_vg = ev_time(lambda: prob.value_grad(x0, U, road, road_idx=ridx))
    t_ro = ev_time(lambda: prob.rollout(x0, U, road, road_idx=ridx))
    fl = FLOP_PER_STAGE_GRAD * N
    out["config5"] = dict(B=B, N=N, value_grad_ms=t_vg, evals_per_s=B / t_vg * 1e3, rollout_ms=t_ro,
                          rollouts_per_s=B / t_ro * 1e3, strict_tflops=B * fl / t_vg / 1e9,
                          roofline_frac=B * fl / t_vg / 1e9 / peak,
                          algorithmic_bytes=B * 696, hbm_gbs=B * 696 / t_vg / 1e6)
    print(json.dumps(out["config5"]), flush=True)
    del prob, x0, U
    # ---- config 3: event-triggered re-solve, N = 40, 2 obstacles, lane potentials; one GPU's share of 64K ----
    B, N = 8192, 40
    lc = wl.lane_change_batch(B, N, 100, seed=3, n_obs=2)
    prob = mpcb200.BatchedMpc(N=N, S=100, model_id=1, road_mode=mpcb200.ROAD_BEZIER,
                              flags=mpcb200.FLAG_LANE | mpcb200.FLAG_OBST, n_obs=2, v_ref=20.0)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    Uc, st_c, x1, Uw, st_w = sc.event_triggered_resolve(prob, lc["x0"], lc["ctrl"], lc["U0"], obs=lc["obs"], seed=3)
    torch.cuda.synchronize(); t_all = time.perf_counter() - t0
    x1d, ctrl, obs = x1, torch.tensor(lc["ctrl"], device=dev), torch.tensor(lc["obs"], device=dev)
    Uw2 = Uc.clone()
    t_warm = ev_time(lambda: prob.solve(x1d, ctrl, Uw2.copy_(Uc), obs=obs), reps=2)
    out["config3"] = dict(B=B, N=N, cold_plus_warm_s=t_all, warm_resolve_ms=t_warm, warm_resolves_per_s=B / t_warm * 1e3,
                          cold_converged=float((st_c.status == 1).double().mean().item()),
                          warm_converged=float((st_w.status == 1).double().mean().item()),
                          cold_inner_mean=float(st_c.inner_iterations.double().mean().item()),
                          warm_inner_mean=float(st_w.inner_iterations.double().mean().item()),
                          profile=prob.last_solve_profile())
    print(json.dumps(out["config3"]), flush=True)
    # ---- config 4: best response, one GPU's share of 8K pairs = 1024 pairs = 2048 vehicles, K = 5 rounds ----
    P, N = 1024, 20
    nv = 2 * P
    lc = wl.lane_change_batch(nv, N, 100, seed=4, n_obs=1)
    x0 = lc["x0"].copy()
    x0[P:, 0] += 30.0; x0[P:, 1] += 3.75; x0[P:, 3] -= 6.0
    prob = mpcb200.BatchedMpc(N=N, S=100, model_id=1, road_mode=mpcb200.ROAD_BEZIER,
                              flags=mpcb200.FLAG_LANE | mpcb200.FLAG_OBST, n_obs=1, v_ref=20.0)
    sc.best_response_rounds(prob, x0, lc["ctrl"], nv, rounds=1)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    r = sc.best_response_rounds(prob, x0, lc["ctrl"], nv, rounds=5)
    torch.cuda.synchronize(); t_br = time.perf_counter() - t0
    out["config4"] = dict(pairs=P, N=N, rounds=5, seconds=t_br, vehicle_solves_per_s=5 * nv / t_br,
                          converged=float((r["status"] == 1).double().mean().item()), delta=r["delta"])
    print(json.dumps(out["config4"]), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "configs.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
