"""GPU-vs-oracle solve parity report (run on the GPU box): distribution of |U_gpu - U_oracle|, psi gaps,
iteration counts.  Writes gpurun_out/parity_report.json.  Test infrastructure (uses the oracle)."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import mpcb200
    from mpcb200 import workloads as wl
    from oracle import oracle as orc
    out = {}
    for N, B in ((12, 512), (20, 512)):
        w = wl.tracking_batch(B, N, 100, seed=40 + N)
        prob = mpcb200.BatchedMpc(N=N, S=100, road_mode=mpcb200.ROAD_TABLE, n_roads=2)
        t0 = time.perf_counter()
        U, _, st = prob.solve(w["x0"], w["road"], w["U0"].copy(), road_idx=w["road_idx"])
        torch.cuda.synchronize()
        t_gpu = time.perf_counter() - t0
        d = orc.default_desc(0, N, 100)
        d.road_mode, d.n_roads = 1, 2
        t0 = time.perf_counter()
        r = orc.solve(d, orc.default_solver_params(N), w["x0"], w["road"], w["U0"], road_idx=w["road_idx"])
        t_cpu = time.perf_counter() - t0
        Ug = U.cpu().numpy()
        err = np.abs(Ug - r["U"]).max(axis=1)
        psi_g, psi_c = st.psi.cpu().numpy(), r["psi"]
        gi, ci = st.inner_iterations.cpu().numpy(), r["inner"]
        q = lambda a, p: float(np.quantile(a, p))
        rep = {
            "B": B, "N": N, "gpu_status": np.bincount(st.status.cpu().numpy(), minlength=7).tolist(),
            "cpu_status": np.bincount(r["status"], minlength=7).tolist(),
            "dU_median": q(err, .5), "dU_p90": q(err, .9), "dU_p99": q(err, .99), "dU_max": float(err.max()),
            "frac_within_1e-5": float((err < 1e-5).mean()), "frac_within_1e-4": float((err < 1e-4).mean()),
            "psi_rel_gap_median": q(np.abs(psi_g - psi_c) / np.abs(psi_c), .5),
            "psi_rel_gap_max": float((np.abs(psi_g - psi_c) / np.abs(psi_c)).max()),
            "gpu_better_or_equal_frac": float((psi_g <= psi_c * (1 + 1e-9)).mean()),
            "same_inner_iters_frac": float((gi == ci).mean()),
            "inner_median_gpu": float(np.median(gi)), "inner_median_cpu": float(np.median(ci)),
            "inner_max_gpu": int(gi.max()), "inner_max_cpu": int(ci.max()),
            "evals_grad_mean_gpu": float(st.evals_grad.double().mean().item()), "evals_grad_mean_cpu": float(r["evals_grad"].mean()),
            "t_gpu_s_first_call": t_gpu, "t_cpu_s": t_cpu, "cpu_threads": orc.num_threads(),
        }
        # where do the outliers differ? (first differing stage input)
        bad = np.where(err >= 1e-5)[0][:5]
        rep["outliers"] = [{"b": int(b), "dU": float(err[b]), "psi_gpu": float(psi_g[b]), "psi_cpu": float(psi_c[b]),
                            "it_gpu": int(gi[b]), "it_cpu": int(ci[b])} for b in bad]
        out[f"N{N}"] = rep
        print(json.dumps(rep))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "parity_report.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
