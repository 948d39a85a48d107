"""Host-side multi-rank logic on CPU: world_size-2 gloo processes (the N > 1 path of bench.py minus the kernels)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, n_total, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from acb200 import sharding
    import bench
    first, count = sharding.shard_range(n_total, rank, world)
    diff, att, phase = bench.make_inputs(count, first)
    # a stand-in for the per-rank accumulator the kernel fills: something only this rank knows
    stats = torch.zeros(16, dtype=torch.float64)
    stats[0] = count; stats[1] = float(diff.astype(np.float64).sum()); stats[15] = float(phase.sum())
    total = sharding.allreduce_episode_stats(stats)
    assert stats[0] == count                          # the local accumulator is left alone
    t = torch.tensor([float(rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)          # the max-over-ranks timing reduction of bench.py
    q.put((rank, first, count, diff, att, phase, total.numpy(), float(t)))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_and_stats_allreduce():
    import torch.multiprocessing as mp
    sys.path.insert(0, ROOT)
    import bench
    world, n_total = 2, 10_001
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_total, q)) for r in range(world)]
    [p.start() for p in procs]
    res = sorted([q.get(timeout=120) for _ in range(world)], key=lambda x: x[0])
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    (r0, f0, c0, d0, a0, p0, tot0, m0), (r1, f1, c1, d1, a1, p1, tot1, m1) = res
    assert (f0, c0, f1, c1) == (0, 5001, 5001, 5000)
    gd, ga, gp = bench.make_inputs(n_total, 0)
    # inputs are a function of the GLOBAL id: the shards concatenate to the single-rank inputs
    assert np.array_equal(np.concatenate([d0, d1]), gd) and np.array_equal(np.concatenate([a0, a1], axis=1), ga)
    assert np.array_equal(np.concatenate([p0, p1]), gp)
    assert np.array_equal(tot0, tot1) and tot0[0] == n_total
    assert abs(tot0[1] - float(gd.astype(np.float64).sum())) < 1e-6 and tot0[15] == float(gp.sum())
    assert m0 == m1 == 2.0


def test_shard_range_properties():
    sys.path.insert(0, ROOT)
    from acb200 import sharding
    for n, w in [(1 << 20, 8), (10, 3), (7, 8), (0, 2)]:
        parts = [sharding.shard_range(n, r, w) for r in range(w)]
        assert parts[0][0] == 0 and sum(c for _, c in parts) == n
        assert all(parts[i][0] + parts[i][1] == parts[i + 1][0] for i in range(w - 1))
        assert max(c for _, c in parts) - min(c for _, c in parts) <= 1
    with pytest.raises(ValueError):
        sharding.shard_range(10, 2, 2)
