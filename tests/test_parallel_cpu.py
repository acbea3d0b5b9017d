"""Host-side multi-process logic (gloo, world_size 2, CPU): sharding and the trajectory all-gather."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, B, q):
    import sys
    sys.path.insert(0, ROOT)
    import mpcb200  # noqa: F401
    from mpcb200 import parallel as par
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        full = torch.arange(B * 3 * 4, dtype=torch.float64).reshape(B, 3, 4)
        mine = par.shard(full)
        lo, hi = par.shard_range(B, rank, world)
        assert mine.shape[0] == hi - lo and torch.equal(mine, full[lo:hi])
        got = par.allgather_trajectories(mine * 1.0, B_total=B)
        q.put((rank, bool(torch.equal(got, full)), tuple(got.shape)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("B", [8, 7])  # equal and ragged shards
def test_shard_and_allgather_world2(B):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, B, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(r[0] for r in res) == [0, 1]
    assert all(r[1] for r in res) and all(r[2] == (B, 3, 4) for r in res)


def test_shard_range_partitions():
    from mpcb200 import parallel as par
    for B in (0, 1, 5, 4096, 65536 + 3):
        for world in (1, 2, 4, 8):
            edges = [par.shard_range(B, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == B
            assert all(edges[i][1] == edges[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in edges]
            assert max(sizes) - min(sizes) <= 1
