#!/usr/bin/env python
"""bench.py — PANOC MPC solves/sec on B200 (BASELINE.json metric), one JSON line on rank 0.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--config metric|2|3|4|5] [--batch B] [--horizon N]

Default (--config metric): step = one pass of the hot path over one batch: cold-start ALM/PANOC solve of `batch`
independent N=20 tracking MPCs per GPU.  The instances follow BASELINE config 2's recipe (Pacejka-6, S=100, 2-road
table, random initial states); the batch is the one BASELINE.json's metric is quoted on ("N=20, batch 64K": 65536 per
GPU, it fits one GPU).  Weak scaling: every rank solves its own `batch` instances, no data-path collective.  Config 2
itself (4096 instances) is measured next to it and reported under "config2_b4096".  The other BASELINE configs are
parity-test cases; --config 2|3|4|5 times them too (same JSON shape, workload named in config.workload) for profiles/.

  value    : solves/s with inputs resident in HBM (one mpcb_solve call per step, CUDA-event timed, max over ranks)
  e2e      : solves/s through the C-ABI host call (numpy buffers -> mpcb_solve_host: pinned staging + H2D + solve +
             D2H inside the timed region); e2e_api: the same through the reference's operator API
             (MPCController.__call__ with (B, 6) states and (B, 2S) centerlines, controller.py:51)
  p50/p99_solve_latency_ms : per-solve latency from the GPU's own clock (every instance stamps %globaltimer when it
             returns its answer); batch_latency_ms is the time of the whole batch
  roofline : FP64-pipe roofline of the dominant kernel (round mode: k_eval_requests, one launch per round): algorithmic
             flops (SURVEY §8d strict count, 3864*N per psi+grad evaluation, 0.455 of that per psi-only evaluation, times
             the kernels' own evaluation counters) / the kernel's launch durations (CUDA events on the launch stream,
             inside the timed region), against the DFMA-chain peak measured in this run (MEASURED_PEAKS.json has no FP64
             entry); useful_frac counts only the evaluations the sequential algorithm (the oracle) needs
  cpu_baseline : the CPU oracle (C++ port of the same spec; alpaqa/CasADi cannot be installed) on a bounded
             sample of the same workload, all host threads.
--impl reference times that oracle as the reference arm.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLOP_PER_STAGE_GRAD = 3864.0   # forward 1758 + reverse 2106 per stage (Pacejka-6, S=100)
VALUE_SHARE = 1758.0 / 3864.0  # a psi-only evaluation
WEIGHTED_FLOP_PER_EVAL_N20 = 234e3  # SURVEY §8d secondary figure (libdevice DFMA-equivalents)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="mpcb200", choices=["mpcb200", "reference"])
    ap.add_argument("--config", default="metric", choices=["metric", "2", "3", "4", "5"],
                    help="metric: BASELINE.json's metric batch (default); 2..5: the other BASELINE configs")
    ap.add_argument("--batch", type=int, default=0, help="instances per GPU (0 = the config's own size)")
    ap.add_argument("--horizon", type=int, default=0, help="0 = the config's own horizon")
    ap.add_argument("--no-api-leg", action="store_true", help="skip the e2e_api leg (MPCController.__call__)")
    ap.add_argument("--cpu-sample", type=int, default=0, help="oracle sample size (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    a = ap.parse_args()
    # sizes of the BASELINE configs: metric 64K x N=20 per GPU; 2: 4096 x N=20; 3: 64K x N=40 over all GPUs;
    # 4: 8192 pairs x N=20 over all GPUs; 5: 1M x N=20 over all GPUs
    if not a.horizon:
        a.horizon = 40 if a.config == "3" else 20
    if not a.batch:
        a.batch = {"metric": 65536, "2": 4096, "3": 65536 // max(a.gpus, 1), "4": 16384 // max(a.gpus, 1),
                   "5": (1 << 20) // max(a.gpus, 1)}[a.config]
    return a


def workload_name(B, N, config="metric"):
    tag = "BASELINE metric batch 64K; config-2 instance recipe" if config == "metric" else "BASELINE config 2"
    return f"{B} independent N={N} Pacejka-6 tracking MPCs per GPU ({tag}), S=100, 2-road table, cold start"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_id):
        self.lines, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(gpu_id), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append((time.perf_counter(), ln.strip()))

    def stop(self, t0=None, t1=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, ln in self.lines:
            if t0 is not None and not (t0 <= ts <= t1 + 0.2):
                continue
            f = [x.strip() for x in ln.split(",")]
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except Exception:
                continue
            for nm, v in zip(names, f[3:7]):
                if v == "Active":
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "power_w_max": float(max(pw)), "samples": len(sm)}


def oracle_setup(N, S=100):
    from oracle import oracle as orc
    d = orc.default_desc(0, N, S)
    d.road_mode, d.n_roads = 1, 2
    return orc, d, orc.default_solver_params(N)


def host_threads():
    """All host cores this process may use (torchrun exports OMP_NUM_THREADS=1: the oracle gets the count explicitly)."""
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def run_oracle(orc, d, sp, w, sample, threads=0):
    threads = threads or host_threads()
    t0 = time.perf_counter()
    r = orc.solve(d, sp, w["x0"][:sample], w["road"], w["U0"][:sample], road_idx=w["road_idx"][:sample],
                  threads=threads, want_times=True)
    return time.perf_counter() - t0, r


def cpu_sample_size(orc, d, sp, w, B, target_s=12.0):
    """Bounded sample: about 10-30 s of CPU work per measurement, scaled from a 64-instance probe."""
    n0 = min(64, B)
    t, _ = run_oracle(orc, d, sp, w, n0)
    per = t / n0
    return int(max(n0, min(B, target_s / max(per, 1e-9))))


def main_reference(args, rank, world):
    if rank != 0:
        return
    from mpcb200 import workloads as wl
    N, B = args.horizon, args.batch
    orc, d, sp = oracle_setup(N)
    w = wl.tracking_batch(B, N, 100, seed=2)
    sample = args.cpu_sample or cpu_sample_size(orc, d, sp, w, B, target_s=6.0)
    for _ in range(min(args.warmup, 1)):
        run_oracle(orc, d, sp, w, min(sample, 64))
    tt = 0.0
    lat = []
    for _ in range(args.steps):
        t, r = run_oracle(orc, d, sp, w, sample)
        tt += t
        lat.append(np.median(r["times"]))
    v = sample * args.steps / tt
    line = {"impl": "reference", "metric": "PANOC MPC solves/sec", "value": v, "unit": "solves/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tt / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(B, N, args.config), "note": "alpaqa/CasADi are not installable here "
                       "(SURVEY §8c): the reference arm is the repo's C++ oracle of the same canonical solver spec"},
            "cpu_baseline": {"value": v, "unit": "solves/s", "cores": host_threads(), "kind": "port",
                             "sample": f"first {sample} instances of the workload per step, OpenMP dynamic over instances"},
            "p50_solve_latency_ms": 1e3 * float(np.median(lat)),
            "e2e": {"value": v, "unit": "solves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main_gpu(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist
    import mpcb200
    from mpcb200 import workloads as wl

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def gather_over_ranks(x):
        if world == 1:
            return [x]
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        out = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return [float(o.item()) for o in out]

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    N, B, S = args.horizon, args.batch, 100
    K, W = args.steps, args.warmup
    w = wl.tracking_batch(B, N, S, seed=2 + rank)  # every rank owns its own instances (weak scaling)
    prob = mpcb200.BatchedMpc(N=N, S=S, road_mode=mpcb200.ROAD_TABLE, n_roads=2)
    x0 = torch.tensor(w["x0"], device=dev)
    road = torch.tensor(w["road"], device=dev)
    ridx = torch.tensor(w["road_idx"], device=dev)
    U0 = torch.tensor(w["U0"], device=dev)
    Ubufs = [U0.clone() for _ in range(K + W)]  # cold start per step, allocated outside the timed region
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    peak_tflops, _ = mpcb200.fp64_peak_tflops()

    # ---- device-resident arm ---------------------------------------------------------------------
    stats = None
    for i in range(W):
        _, _, stats = prob.solve(x0, road, Ubufs[i], road_idx=ridx)
    barrier()
    gpu_id = local_rank
    try:
        gpu_id = "GPU-" + str(torch.cuda.get_device_properties(local_rank).uuid)
    except Exception:
        pass
    sampler = ClockSampler(gpu_id)
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    l0 = prob.launch_count()
    barrier()
    t_wall0 = time.perf_counter()
    all_stats = []
    profs = []
    for i in range(K):
        flush.zero_()  # L2 flush between timed iterations (not timed)
        evs[i][0].record()
        _, _, st = prob.solve(x0, road, Ubufs[W + i], road_idx=ridx)
        evs[i][1].record()
        all_stats.append(st)
        profs.append(prob.last_solve_profile())
    lat_ms = prob.last_solve_latency_ms(B).cpu().numpy()   # per-solve latency of the last timed step (GPU clock stamps)
    barrier()
    t_wall1 = time.perf_counter()
    launches = prob.launch_count() - l0
    step_ms = [a.elapsed_time(b) for a, b in evs]
    clocks = sampler.stop(t_wall0, t_wall1)
    t_dev = max_over_ranks(sum(step_ms) * 1e-3)
    value = world * B * K / t_dev

    st = all_stats[-1]
    status = st.status.cpu().numpy()
    n_val = float(st.evals_value.sum().item())
    n_grad = float(st.evals_grad.sum().item())
    flop_eval = FLOP_PER_STAGE_GRAD * N
    alg_flops = (n_grad + VALUE_SHARE * n_val) * flop_eval
    prof = profs[-1]
    rounds_mode = prof["mode"] == "rounds"
    # dominant kernel: the evaluation launches in round mode (their summed CUDA-event durations), else the one kernel
    k_ms = float(prof["eval_ms"]) if rounds_mode else float(step_ms[-1])
    n_k_launches = prof["eval_launches"]
    achieved = alg_flops / (k_ms * 1e-3) / 1e12
    weighted = (n_grad + VALUE_SHARE * n_val) * WEIGHTED_FLOP_PER_EVAL_N20 * (N / 20.0) / (k_ms * 1e-3) / 1e12
    alg_bytes = B * (6 * 8 + 2 * N * 8 * 2 + 4 + 4 * 4 + 2 * 8 + 4 * 4)  # x0, U in/out, road_idx, status/iters/evals, psi/res
    roofline = {"bound": "fp64", "kernel": "k_eval_requests<Pacejka6>" if rounds_mode else "k_solve<Pacejka6>",
                "achieved": achieved, "peak": peak_tflops,
                "unit": "TFLOP/s", "frac": achieved / peak_tflops, "traffic": _ncu_traffic(rounds_mode),
                "launches_per_step": n_k_launches, "avg_launch_ms": k_ms / max(n_k_launches, 1),
                "flop_per_launch": alg_flops / max(n_k_launches, 1),
                "bulk": _bulk_roofline(prof, flop_eval, peak_tflops),
                "kernel_share_of_step": k_ms / float(step_ms[-1]),
                "solve_profile": prof,
                "peak_source": "DFMA-chain probe measured in this run (mpcb_fp64_peak); MEASURED_PEAKS.json has no FP64 figure",
                "flop_per_eval": flop_eval, "evals_grad_per_step": n_grad, "evals_value_per_step": n_val,
                "weighted_achieved": weighted, "weighted_frac": weighted / peak_tflops,
                "hbm": {"algorithmic_bytes_per_step": alg_bytes, "achieved_gbs": alg_bytes / (k_ms * 1e-3) / 1e9,
                        "peak_gbs": _hbm_peak()}}

    # ---- cross-check of the kernel durations with CUDA events: one untimed step through the host-driven loop -------
    # (the default loop is a CUDA graph WHILE node: its per-launch durations are %globaltimer stamps written by the kernels,
    # events cannot bracket nodes inside the loop body; MPCB_LOOP=host launches the same kernels from the host with
    # cudaEventRecord around every launch on the launch stream)
    os.environ["MPCB_LOOP"] = "host"
    try:
        prob_h = mpcb200.BatchedMpc(N=N, S=S, road_mode=mpcb200.ROAD_TABLE, n_roads=2)
    finally:
        os.environ.pop("MPCB_LOOP", None)
    Uh0 = U0.clone()
    prob_h.solve(x0, road, U0.clone(), road_idx=ridx)
    a_ev, b_ev = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a_ev.record()
    prob_h.solve(x0, road, Uh0, road_idx=ridx)
    b_ev.record()
    torch.cuda.synchronize()
    ph = prob_h.last_solve_profile()
    roofline["launch_timing"] = ("graph-driven loop: per-launch durations from %globaltimer stamps at the start of each evaluation "
                                 "launch and at the end of each round (trans_ms includes node scheduling)")
    roofline["cuda_event_cross_check"] = {"loop": "host-driven (MPCB_LOOP=host), one untimed step", "step_ms": a_ev.elapsed_time(b_ev),
                                          "eval_ms": ph["eval_ms"], "trans_ms": ph["trans_ms"], "rounds": ph["rounds"],
                                          "achieved": alg_flops / (ph["eval_ms"] * 1e-3) / 1e12 if rounds_mode else None,
                                          "bit_identical_U": bool(torch.equal(Uh0, Ubufs[W + K - 1]))}
    del prob_h

    # ---- end-to-end arm: host buffers through mpcb_solve_host ---------------------------------------
    Uh = [w["U0"].copy() for _ in range(K + 1)]
    prob.solve_host(w["x0"], w["road"], Uh[K], road_idx=w["road_idx"])  # warm (allocates pinned staging)
    barrier()
    t0 = time.perf_counter()
    for i in range(K):
        prob.solve_host(w["x0"], w["road"], Uh[i], road_idx=w["road_idx"])
    t_e2e = time.perf_counter() - t0
    barrier()
    t_e2e = max_over_ranks(t_e2e)
    h2d, d2h = prob.last_transfer_bytes()

    # ---- end-to-end through the reference's operator API: MPCController.__call__ (controller.py:51) ------------
    # (B, 6) states and (B, 2S) centerlines as a reference user holds them; the call dedupes the centerlines into a road
    # table, stages, copies, solves and copies back; cold start per step like the other arms
    e2e_api = None
    if not args.no_api_leg:
        import warnings
        from mpcb200 import main as ref_main
        from mpcb200.car_dynamics import KinematicBicyclePacejka
        from mpcb200.controller import MPCController
        model = KinematicBicyclePacejka()
        aprob = ref_main.create_casadi_problem(model, N, S, 1.0, 1.0, 0.32)
        aprob.param = np.concatenate((w["x0"][0], w["road"][0], ref_main.VEHICLE_PARAM))
        ctl = MPCController(model, aprob, N)
        centerlines = np.ascontiguousarray(w["road"][w["road_idx"]])
        Ka = max(1, min(K, 3))
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            ctl(w["x0"], centerlines)   # warm (creates the table-road handle, allocates pinned staging)
            barrier()
            t0 = time.perf_counter()
            for i in range(Ka):
                ctl.U = np.tile([1.0, 0.0], (B, N))   # controller.py:20 cold start
                ctl.λ = np.zeros((B, 6 * N))
                Ua = ctl(w["x0"], centerlines)
            t_api = time.perf_counter() - t0
        barrier()
        t_api = max_over_ranks(t_api)
        api_mpc = aprob.mpc(mpcb200.ROAD_TABLE, 2)
        ah2d, ad2h = api_mpc.last_transfer_bytes()
        e2e_api = {"value": world * B * Ka / t_api, "unit": "solves/s", "steps": Ka, "h2d_bytes_per_step": ah2d,
                   "d2h_bytes_per_step": ad2h, "road_layout": ctl.last_stats["road_layout"],
                   "host_input_bytes_per_step": int(w["x0"].nbytes + centerlines.nbytes + 2 * B * 2 * N * 8),
                   "failures": int(ctl.failures), "call": "MPCController((B,6) states, (B,2S) centerlines) -> (B,2N)"}

    # ---- single-instance latency (the reference's deployment shape: one vehicle, one tick) ------------
    lat1 = []
    U1 = [U0[:1].clone() for _ in range(12)]
    for i in range(12):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        prob.solve(x0[:1], road, U1[i], road_idx=ridx[:1])
        b.record()
        torch.cuda.synchronize()
        lat1.append(a.elapsed_time(b))
    conv = sum_over_ranks(float((status == 1).sum())) / (world * B)
    rounds_all = gather_over_ranks(float(prof["rounds"]))
    step_ms_all = gather_over_ranks(float(np.mean(step_ms)))

    # ---- BASELINE config 2 itself: 4096 instances on one GPU (reported next to the headline) ----------------
    cfg2 = None
    if B != 4096 and args.config == "metric":
        B2 = 4096
        w2 = wl.tracking_batch(B2, N, S, seed=2 + rank)
        x2, r2 = torch.tensor(w2["x0"], device=dev), torch.tensor(w2["road_idx"], device=dev)
        U2 = [torch.tensor(w2["U0"], device=dev) for _ in range(3)]
        t2 = []
        for i in range(3):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            _, _, s2 = prob.solve(x2, road, U2[i], road_idx=r2)
            b.record()
            torch.cuda.synchronize()
            t2.append(a.elapsed_time(b))
        launches += 0  # outside the timed region
        cfg2 = {"batch": B2, "ms_per_batch": float(np.median(t2[1:])), "solves_per_s": B2 / float(np.median(t2[1:])) * 1e3,
                "converged_frac": float((s2.status == 1).double().mean().item()), "solve_profile": prob.last_solve_profile()}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        orc, d, sp = oracle_setup(N)
        sample = args.cpu_sample or cpu_sample_size(orc, d, sp, w, B)
        t, r = run_oracle(orc, d, sp, w, sample)
        cpu = {"value": sample / t, "unit": "solves/s", "cores": host_threads(), "kind": "port",
               "sample": f"first {sample} instances of the same workload, OpenMP over instances",
               "p50_solve_latency_ms": 1e3 * float(np.median(r["times"])),
               "evals_grad_per_solve": float(r["evals_grad"].mean()), "evals_value_per_solve": float(r["evals_value"].mean()),
               "build": "oracle/Makefile: g++ -O2 -ffp-contract=off -fopenmp (measured equal to -O3 -march=native within 5 %)"}
        # the sequential algorithm's evaluations (what the oracle needs for the same solves) as the useful work
        useful_units = B * (cpu["evals_grad_per_solve"] + VALUE_SHARE * cpu["evals_value_per_solve"])
        useful = useful_units * flop_eval / (k_ms * 1e-3) / 1e12
        roofline["useful_achieved"] = useful
        roofline["useful_frac"] = useful / peak_tflops
        roofline["useful_note"] = ("oracle-equivalent evaluations (sequential back-tracking, no speculation) x flop_per_eval / "
                                   "kernel time / peak; GPU evaluations per solve / oracle's = "
                                   f"{(n_grad + VALUE_SHARE * n_val) / max(useful_units, 1):.2f}")

    if rank == 0:
        line = {"metric": "PANOC MPC solves/sec", "value": value, "unit": "solves/s", "n_gpus": world, "steps": K,
                "warmup": W, "ms_per_step": 1e3 * t_dev / K, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": workload_name(B, N, args.config), "batch_per_gpu": B, "horizon": N,
                           "l2": "flushed between timed steps (256 MiB memset, untimed)",
                           "timing": "CUDA events per step on the launch stream, summed, max over ranks; e2e: host wall clock around the synchronous host call"},
                "p50_solve_latency_ms": float(np.median(lat_ms)), "p99_solve_latency_ms": float(np.quantile(lat_ms, 0.99)),
                "max_solve_latency_ms": float(lat_ms.max()), "batch_latency_ms": float(np.median(step_ms)),
                "latency_note": "per-solve: GPU %globaltimer stamp when the instance returned its answer, minus the start of the "
                                "batch on the GPU (last timed step, rank 0); batch_latency_ms: the whole tick",
                "single_instance_latency_ms_p50": float(np.median(lat1[2:])),
                "rounds_per_rank": [int(x) for x in rounds_all], "step_ms_per_rank": step_ms_all,
                "converged_frac": conv, "inner_iterations_mean": float(st.inner_iterations.double().mean().item()),
                "e2e": {"value": world * B * K / t_e2e, "unit": "solves/s", "h2d_bytes_per_step": h2d,
                        "d2h_bytes_per_step": d2h},
                "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline,
                "solver_note": "hessian_step_size_heuristic off (ignored: alpaqa 0.0.1's rule is not recoverable), both arms"}
        if e2e_api:
            line["e2e_api"] = e2e_api
        if cfg2:
            line["config2_b4096"] = cfg2
        if cpu:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def _bulk_roofline(prof, flop_eval, peak):
    """The same roofline over the throughput-bound rounds only (more requests than the evaluation launch has lanes): the
    step-wide figure averages in ~2 000 launches that serve a handful of slow instances at the latency floor."""
    b = prof.get("bulk")
    if not b or not b["rounds"] or b["eval_ms"] <= 0:
        return None
    fl = (b["grad_requests"] + VALUE_SHARE * b["value_requests"]) * flop_eval
    ach = fl / (b["eval_ms"] * 1e-3) / 1e12
    return {"rounds": b["rounds"], "eval_ms": b["eval_ms"], "trans_ms": b["trans_ms"], "achieved": ach, "frac": ach / peak,
            "avg_launch_ms": b["eval_ms"] / b["rounds"], "evaluations_per_s": (b["grad_requests"] + b["value_requests"]) / (b["eval_ms"] * 1e-3)}


def _ncu_traffic(rounds_mode):
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel from the committed ncu capture (one bulk-phase
    launch; the profiles/rNN_ncu_traffic.json file says which), or None."""
    if not rounds_mode:
        return None
    for name in ("r02_ncu_traffic.json", "r01_ncu_traffic.json"):
        try:
            t = json.load(open(os.path.join(ROOT, "profiles", name)))
            return {"dram_bytes_per_launch": t["dram_bytes_per_launch"], "of": t["capture"], "note": t["note"], "file": name}
        except Exception:
            continue
    return None


class _Dist:
    """One process per GPU: NCCL group, barrier + max-over-ranks timing as the bench contract asks."""

    def __init__(self, local_rank, world):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.world = torch, dist, world
        torch.cuda.set_device(local_rank)
        self.dev = torch.device("cuda", local_rank)
        if world > 1:
            dist.init_process_group("nccl", device_id=self.dev)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, x, op="max"):
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX if op == "max" else self.dist.ReduceOp.SUM)
        return float(t.item())

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def _timed_steps(D, K, W, step, sampler_gpu):
    """W warm-up steps, then K timed steps bracketed by barrier + synchronize, CUDA events per step, max over ranks."""
    torch = D.torch
    for i in range(W):
        step(i)
    D.barrier()
    sampler = ClockSampler(sampler_gpu)
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=D.dev)
    D.barrier()
    t0 = time.perf_counter()
    for i in range(K):
        flush.zero_()
        evs[i][0].record()
        step(W + i)
        evs[i][1].record()
    D.barrier()
    t1 = time.perf_counter()
    ms = [a.elapsed_time(b) for a, b in evs]
    return D.reduce(sum(ms) * 1e-3), ms, sampler.stop(t0, t1)


def _gpu_id(torch, local_rank):
    try:
        return "GPU-" + str(torch.cuda.get_device_properties(local_rank).uuid)
    except Exception:
        return local_rank


def _line(args, world, value, unit, t_dev, scaling, workload, extra):
    K = args.steps
    line = {"metric": "PANOC MPC solves/sec" if unit == "solves/s" else "psi + grad psi evaluations/sec", "value": value,
            "unit": unit, "n_gpus": world, "steps": K, "warmup": args.warmup, "ms_per_step": 1e3 * t_dev / K,
            "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload, "baseline_config": args.config,
                       "l2": "flushed between timed steps (256 MiB memset, untimed)",
                       "timing": "CUDA events per step on the launch stream, summed, max over ranks"}}
    line.update(extra)
    return line


def main_config3(args, rank, local_rank, world):
    """BASELINE config 3: event-triggered re-solve sweep — 64K Kinematic-4 lane-change instances over all GPUs, N = 40,
    2 moving obstacles (dnf) + lane potentials; step = the warm re-solve of every local instance from a state
    perturbed by N(0, 0.02^2) (the cold solves that provide the warm starts run once, untimed)."""
    import mpcb200
    from mpcb200 import scenarios as sc, workloads as wl
    D = _Dist(local_rank, world)
    torch = D.torch
    B, N = args.batch, args.horizon
    lc = wl.lane_change_batch(B, N, 100, seed=3 + rank, n_obs=2)
    prob = mpcb200.BatchedMpc(N=N, S=100, model_id=1, road_mode=mpcb200.ROAD_BEZIER,
                              flags=mpcb200.FLAG_LANE | mpcb200.FLAG_OBST, n_obs=2, v_ref=20.0)
    Uc, st_c, x1, Uw, st_w = sc.event_triggered_resolve(prob, lc["x0"], lc["ctrl"], lc["U0"], obs=lc["obs"], seed=3 + rank)
    ctrl, obs = torch.tensor(lc["ctrl"], device=D.dev), torch.tensor(lc["obs"], device=D.dev)
    Ub = [Uc.clone() for _ in range(args.steps + args.warmup)]
    stats = {}

    def step(i):
        stats["st"] = prob.solve(x1, ctrl, Ub[i], obs=obs)[2]

    l0 = prob.launch_count()
    t_dev, ms, clocks = _timed_steps(D, args.steps, args.warmup, step, _gpu_id(torch, local_rank))
    launches = (prob.launch_count() - l0) * args.steps // (args.steps + args.warmup)
    st = stats["st"]
    lat = prob.last_solve_latency_ms(B).cpu().numpy()
    conv = D.reduce(float((st.status == 1).sum().item()), "sum") / (world * B)
    if rank == 0:
        print(json.dumps(_line(args, world, world * B * args.steps / t_dev, "solves/s", t_dev, "strong",
                               f"{world * B} Kinematic-4 lane-change MPCs over {world} GPU(s) ({B} per GPU), N={N}, 2 moving "
                               "obstacles + lane potentials, Bezier roads; warm re-solve after a N(0, 0.02^2) state perturbation",
                               {"converged_frac": conv, "cold_inner_mean": float(st_c.inner_iterations.double().mean().item()),
                                "warm_inner_mean": float(st.inner_iterations.double().mean().item()),
                                "p50_solve_latency_ms": float(np.median(lat)), "p99_solve_latency_ms": float(np.quantile(lat, .99)),
                                "batch_latency_ms": float(np.median(ms)), "gpu_launches": int(launches), "clocks": clocks,
                                "solve_profile": prob.last_solve_profile()})), flush=True)
    D.close()


def main_config4(args, rank, local_rank, world):
    """BASELINE config 4: two-vehicle iterated best response — 8192 scenario pairs (16384 vehicles) over all GPUs, K = 5
    rounds; every round: payoff-based lane decision (game_theory.py) -> Bezier path -> solve the local vehicles -> roll
    out -> NCCL all-gather of the [B_local, N, 4] trajectories -> rebuild the opponents' obstacle records."""
    import mpcb200
    from mpcb200 import parallel, scenarios as sc, workloads as wl
    D = _Dist(local_rank, world)
    torch = D.torch
    N = args.horizon
    nv = args.batch * world
    w4 = wl.game_pairs(nv // 2, seed=4)
    lo, hi = parallel.shard_range(nv, rank, world)
    prob = mpcb200.BatchedMpc(N=N, S=100, model_id=1, road_mode=mpcb200.ROAD_BEZIER,
                              flags=mpcb200.FLAG_LANE | mpcb200.FLAG_OBST, n_obs=1, v_ref=20.0)
    res = {}

    def step(i):
        res["r"] = sc.best_response_rounds(prob, w4["x0"][lo:hi], None, nv, rounds=5, cars=w4, timing=True)

    l0 = prob.launch_count()
    t_dev, ms, clocks = _timed_steps(D, args.steps, args.warmup, step, _gpu_id(torch, local_rank))
    launches = (prob.launch_count() - l0) * args.steps // (args.steps + args.warmup)
    r = res["r"]
    conv = D.reduce(float((r["status"] == 1).sum().item()), "sum") / (5 * nv)
    if rank == 0:
        print(json.dumps(_line(args, world, 5 * nv * args.steps / t_dev, "solves/s", t_dev, "strong",
                               f"{nv // 2} two-vehicle best-response scenarios ({nv} Kinematic-4 MPCs, {hi - lo} per GPU), N={N}, "
                               "5 rounds of payoff lane decision -> Bezier path -> solve -> rollout -> all-gather of trajectories",
                               {"converged_frac": conv, "rounds": 5, "delta": r["delta"], "lane_changes_frac": r.get("lane_change_frac"),
                                "allgather_ms_per_round": r.get("gather_ms"), "solve_ms_per_round": r.get("solve_ms"),
                                "allgather_bytes_per_rank_per_round": int((hi - lo) * N * 4 * 8),
                                "collective": "torch.distributed all_gather_into_tensor (NCCL)" if world > 1 else "none (1 GPU)",
                                "gpu_launches": int(launches), "clocks": clocks})), flush=True)
    D.close()


def main_config5(args, rank, local_rank, world):
    """BASELINE config 5: 1M-instance open-loop rollout + gradient sweep over all GPUs (the roofline anchor): step = one
    mpcb_value_grad call over the local shard."""
    import mpcb200
    from mpcb200 import workloads as wl
    D = _Dist(local_rank, world)
    torch = D.torch
    B, N = args.batch, args.horizon
    w = wl.tracking_batch(B, N, 100, seed=5 + rank, random_U=True)
    prob = mpcb200.BatchedMpc(N=N, S=100, road_mode=mpcb200.ROAD_TABLE, n_roads=2)
    x0, road = torch.tensor(w["x0"], device=D.dev), torch.tensor(w["road"], device=D.dev)
    ridx, U = torch.tensor(w["road_idx"], device=D.dev), torch.tensor(w["U0"], device=D.dev)
    out = (torch.empty(B, dtype=torch.float64, device=D.dev), torch.empty(B, 2 * N, dtype=torch.float64, device=D.dev))
    peak, _ = mpcb200.fp64_peak_tflops()

    def step(i):
        prob.value_grad(x0, U, road, road_idx=ridx, out=out)

    t_dev, ms, clocks = _timed_steps(D, args.steps, args.warmup, step, _gpu_id(torch, local_rank))
    flops = FLOP_PER_STAGE_GRAD * N * B * args.steps
    ach = flops / (sum(ms) * 1e-3) / 1e12
    if rank == 0:
        print(json.dumps(_line(args, world, world * B * args.steps / t_dev, "evaluations/s", t_dev, "strong",
                               f"{world * B}-instance open-loop rollout + gradient sweep over {world} GPU(s) ({B} per GPU), Pacejka-6, "
                               f"N={N}, random inputs in the box",
                               {"gpu_launches": args.steps, "clocks": clocks,
                                "roofline": {"bound": "fp64", "kernel": "k_value_grad<Pacejka6>", "achieved": ach, "peak": peak,
                                             "unit": "TFLOP/s", "frac": ach / peak, "traffic": None,
                                             "flop_per_eval": FLOP_PER_STAGE_GRAD * N,
                                             "hbm": {"algorithmic_bytes_per_step": B * 696,
                                                     "achieved_gbs": B * 696 * args.steps / (sum(ms) * 1e-3) / 1e9,
                                                     "peak_gbs": _hbm_peak()}}})), flush=True)
    D.close()


def _hbm_peak():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        return 6650.0  # B200_PROFILING.md fallback


if __name__ == "__main__":
    a = parse()
    rank = int(os.environ.get("RANK", 0))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if a.impl == "reference":
        if a.config in ("metric", "2"):
            main_reference(a, rank, world)
        elif rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": f"the reference arm times the metric batch and config 2 only (asked: config {a.config})"}))
    elif a.config == "3":
        main_config3(a, rank, local_rank, world)
    elif a.config == "4":
        main_config4(a, rank, local_rank, world)
    elif a.config == "5":
        main_config5(a, rank, local_rank, world)
    else:
        main_gpu(a, rank, local_rank, world)
