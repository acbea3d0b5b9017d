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


class _StubProb:
    """Stand-in for BatchedMpc in the host-logic tests (CPU, no kernels): `solve` moves U towards a target that
    depends on the obstacle record, `rollout` integrates x along U.  Only the sequencing is under test."""
    nx, N, n, m = 4, 5, 10, 20

    class desc:
        Ts, flags = 0.05, 0

    device = torch.device("cpu")

    def _t(self, x, dtype=torch.float64):
        return torch.as_tensor(x, dtype=dtype).contiguous()

    def solve(self, x0, road, U, obs=None, sp=None, **kw):
        from types import SimpleNamespace
        U = 0.5 * U + 0.1 * obs[:, 0, 0:1] - 0.01 * obs[:, 0, 4:5] + 0.001 * road[:, 2:3]
        return U, None, SimpleNamespace(status=torch.ones(x0.shape[0], dtype=torch.int32))

    def rollout(self, x0, U, road, obs=None, **kw):
        steps = U.view(-1, self.N, 2).cumsum(dim=1)
        traj = x0[:, None, :].repeat(1, self.N, 1)
        traj[:, :, 0] += steps[:, :, 0]
        traj[:, :, 3] += steps[:, :, 1]
        return traj, None, None


def _br_inputs(nv):
    g = torch.Generator().manual_seed(3)
    x0 = torch.rand(nv, 4, dtype=torch.float64, generator=g)
    ctrl = torch.rand(nv, 12, dtype=torch.float64, generator=g)
    return x0, ctrl


def _br_worker(rank, world, port, nv, q):
    import sys
    sys.path.insert(0, ROOT)
    import mpcb200  # noqa: F401
    from mpcb200 import parallel as par, scenarios as sc
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        x0, ctrl = _br_inputs(nv)
        r = sc.best_response_rounds(_StubProb(), par.shard(x0), par.shard(ctrl), nv, rounds=3)
        q.put((rank, r["traj"].numpy(), r["U"].numpy(), r["delta"]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("nv", [8, 6])  # 6 vehicles over 2 ranks: opponents live on the other rank either way
def test_best_response_rounds_world2_equals_world1(nv):
    """Config 4 sequencing: sharded over 2 ranks with the all-gather == unsharded."""
    import sys
    sys.path.insert(0, ROOT)
    import mpcb200  # noqa: F401
    from mpcb200 import parallel as par, scenarios as sc
    x0, ctrl = _br_inputs(nv)
    ref = sc.best_response_rounds(_StubProb(), x0, ctrl, nv, rounds=3, rank=0, world=1)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_br_worker, args=(r, 2, port, nv, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, traj, U, delta in res:
        np.testing.assert_array_equal(traj, ref["traj"].numpy())
        lo, hi = par.shard_range(nv, rank, 2)
        np.testing.assert_array_equal(U, ref["U"].numpy()[lo:hi])
        assert delta[1:] == ref["delta"][1:]


def test_scenario_helpers():
    import sys
    sys.path.insert(0, ROOT)
    import mpcb200  # noqa: F401
    from mpcb200 import scenarios as sc
    # event_triggered.py:31-37: ego = Car(0, 0, 10, 0, 0, 0), t0 = 1, t1 = 2 -> 30
    assert sc.predict_position(0.0, 10.0, 0.0, 1, 2) == 30.0
    assert sc.predict_position(1.0, 2.0, 4.0, 0.5, 0.5) == 1.0 + 2.0 + 2.0
    U = torch.arange(12, dtype=torch.float64).reshape(2, 6)
    assert sc.shift_warm_start(U, 3).tolist() == [[2, 3, 4, 5, 4, 5], [8, 9, 10, 11, 10, 11]]
    assert sc.opponent_index(6).tolist() == [3, 4, 5, 0, 1, 2]
    x = torch.tensor([[1.0, 2.0, 0.1, 10.0]])
    traj = x[:, None, :].repeat(1, 4, 1)
    traj[0, -1, 3] = 12.0
    o = sc.obstacles_from_trajectories(x, traj, 0.05)
    assert o.shape == (1, 1, 5)
    assert o[0, 0].tolist() == pytest.approx([1.0, 2.0, 0.1, 10.0, 2.0 / 0.2])
