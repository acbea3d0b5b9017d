"""Config 4 under NCCL: shard invariance and the cost of the one collective.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/config4_shard_check.py

Every rank solves its contiguous shard of the 2P vehicles for K rounds with the NCCL all-gather of [B_local, N, 4]
trajectories; rank 0 also runs the whole problem alone (no collective) and every rank compares its shard bit for bit.
Prints one JSON line (rank 0) with the per-round solve / all-gather times."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    import mpcb200
    from mpcb200 import parallel, scenarios as sc, workloads as wl
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    P = int(os.environ.get("PAIRS", "8192"))
    N, K = 20, 5
    nv = 2 * P
    w4 = wl.game_pairs(P, seed=4)
    prob = mpcb200.BatchedMpc(N=N, S=100, model_id=1, road_mode=mpcb200.ROAD_BEZIER,
                              flags=mpcb200.FLAG_LANE | mpcb200.FLAG_OBST, n_obs=2, v_ref=20.0)
    lo, hi = parallel.shard_range(nv, rank, world)
    sc.best_response_rounds(prob, w4["x0"][lo:hi], None, nv, rounds=1, cars=w4)   # warm
    r = sc.best_response_rounds(prob, w4["x0"][lo:hi], None, nv, rounds=K, cars=w4, timing=True)
    # the whole problem on this GPU alone: the reference bits.  (The process group stays initialised, so the exchange is
    # bypassed explicitly: at world = 1 it is the identity.)
    ref = r
    if world > 1:
        saved = parallel.allgather_trajectories
        parallel.allgather_trajectories = lambda local_t, B_total=None, group=None: local_t
        try:
            ref = sc.best_response_rounds(prob, w4["x0"], None, nv, rounds=K, cars=w4, rank=0, world=1)
        finally:
            parallel.allgather_trajectories = saved
    same_U = bool(torch.equal(r["U"], ref["U"][lo:hi]))
    same_traj = bool(torch.equal(r["traj"], ref["traj"]))
    same_target = bool(torch.equal(r["target"], ref["target"][:, lo:hi]))
    ok = torch.tensor([int(same_U and same_traj and same_target)], device=dev)
    if world > 1:
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(json.dumps({"config": 4, "pairs": P, "vehicles": nv, "world": world, "rounds": K,
                          "shard_invariant_bits": bool(ok.item()), "converged": float((r["status"] == 1).double().mean().item()),
                          "solve_ms_per_round": r["solve_ms"], "allgather_ms_per_round": r["gather_ms"],
                          "allgather_bytes_per_rank": (hi - lo) * N * 4 * 8, "delta": r["delta"],
                          "lane_change_frac": r["lane_change_frac"]}), flush=True)
    if world > 1:
        dist.destroy_process_group()
    sys.exit(0 if ok.item() else 1)


if __name__ == "__main__":
    main()
