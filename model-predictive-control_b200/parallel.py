"""Multi-GPU plumbing for the hot path: one process per GPU (torch.distributed; NCCL on GPUs, gloo on CPU).

MPC instances are independent, so the path shards with no data-path collective:
`shard_range` gives every rank a contiguous block of instances and the rank solves it with its own handle.
The one exchange step in the north star is the game-theory configuration (config 4): after every
best-response round each rank's predicted trajectories `[B_local, N, 4]` are all-gathered so that every
vehicle's solve can treat the other vehicle's trajectory as a moving obstacle (`allgather_trajectories`).
Nothing here touches the oracle or a CPU compute path.
"""
import torch
import torch.distributed as dist


def shard_range(B, rank, world):
    """Contiguous block [lo, hi) of rank `rank` out of B instances: sizes differ by at most one, rank order
    preserved (concatenating the shards in rank order restores the batch)."""
    base, rem = divmod(int(B), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard(t, rank=None, world=None):
    """This rank's rows of a batch-major tensor / array."""
    rank = dist.get_rank() if rank is None else rank
    world = dist.get_world_size() if world is None else world
    lo, hi = shard_range(t.shape[0], rank, world)
    return t[lo:hi]


def allgather_trajectories(local, B_total=None, group=None):
    """All-gather of batch-major tensors with (possibly) ragged shard sizes.

    local: [B_local, ...] on this rank's device.  Returns [B_total, ...] in rank order on every rank.
    Equal shards use one all_gather_into_tensor on the caller's stream (NCCL: straight from the buffer the
    solver wrote); ragged shards are padded to the largest shard first."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local
    world = dist.get_world_size(group)
    n_local = torch.tensor([local.shape[0]], dtype=torch.int64, device=local.device)
    sizes = [torch.zeros_like(n_local) for _ in range(world)]
    dist.all_gather(sizes, n_local, group=group)
    sizes = [int(s.item()) for s in sizes]
    mx = max(sizes)
    local = local.contiguous()
    if all(s == mx for s in sizes):
        out = torch.empty((world * mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(out, local, group=group)
    else:
        pad = torch.zeros((mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        pad[:local.shape[0]] = local
        buf = torch.empty((world * mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(buf, pad, group=group)
        out = torch.cat([buf[r * mx:r * mx + sizes[r]] for r in range(world)], dim=0)
    if B_total is not None and out.shape[0] != B_total:
        raise ValueError(f"gathered {out.shape[0]} rows, expected {B_total}")
    return out
