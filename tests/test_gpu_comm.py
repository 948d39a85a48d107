"""acb200_comm_* / acb200_episode_stats: the statistics all-reduce through the C ABI (NCCL resolved with dlopen).  One GPU is
enough to drive the whole code path: a communicator of world size 1 is a real ncclCommInitRank + ncclAllReduce.  The
two-rank sum is checked by tools/probes/comm_check.py under torchrun (profiles/r02_comm_check_2gpu.log)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def test_episode_stats_copy_and_world_of_one_allreduce():
    import acb200
    ctx = acb200.Context(0)
    lib, h = ctx.lib, ctx.handle
    stats = torch.arange(16, device="cuda", dtype=torch.float64) * 1.5 + 0.25
    out = torch.zeros_like(stats)
    p = lambda t: C.c_void_p(t.data_ptr())
    # do_allreduce = 0: a copy, no communicator needed
    assert lib.acb200_episode_stats(h, p(stats), p(out), 16, 0) == 0
    torch.cuda.synchronize()
    assert torch.equal(out, stats)
    # do_allreduce = 1 without a communicator: refused, not silently local
    assert lib.acb200_episode_stats(h, p(stats), p(out), 16, 1) == acb200._cabi.ENOSUP
    assert b"communicator" in lib.acb200_last_error(h)
    assert lib.acb200_episode_stats(h, p(stats), p(out), 0, 0) == acb200._cabi.EVALUE
    assert lib.acb200_episode_stats(h, None, p(out), 16, 0) == acb200._cabi.ENULL
    # a communicator of one rank: the real NCCL calls
    uid = (C.c_ubyte * 128)()
    rc = lib.acb200_comm_unique_id(h, uid)
    if rc == acb200._cabi.ENOSUP:
        pytest.skip("no NCCL library in this process: " + lib.acb200_last_error(h).decode())
    assert rc == 0 and any(uid)
    assert lib.acb200_comm_init(h, 2, 5, uid) == acb200._cabi.EVALUE            # rank outside the world
    assert lib.acb200_comm_init(h, 1, 0, uid) == 0, lib.acb200_last_error(h)
    assert lib.acb200_comm_init(h, 1, 0, uid) == acb200._cabi.EVALUE            # already has one
    info = acb200.sharding.comm_info(ctx)
    assert info["world"] == 1 and info["rank"] == 0 and info["nccl_version"] >= 20000
    out.zero_()
    l0 = ctx.launch_count
    assert lib.acb200_episode_stats(h, p(stats), p(out), 16, 1) == 0, lib.acb200_last_error(h)
    torch.cuda.synchronize()
    assert torch.equal(out, stats) and ctx.launch_count == l0 + 1
    # through the Python helper, in place and on a side stream
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        got = acb200.sharding.allreduce_episode_stats(stats, ctx=ctx)
    s.synchronize()
    assert torch.equal(got, stats)
    assert lib.acb200_comm_destroy(h) == 0 and acb200.sharding.comm_info(ctx)["world"] == 0
    assert lib.acb200_comm_destroy(h) == 0                                      # idempotent


def test_env_statistics_through_the_c_abi_match_the_env_accumulator():
    import acb200
    env = acb200.BatchedRateControlEnv(4096, controller="rate_pid", precision=32, seed=3, episode_length=0.4)
    env.reset()
    for _ in range(30):
        env.step()
    assert acb200.sharding.init_comm(env.ctx) is False                          # one rank, no process group: nothing to set up
    got = acb200.sharding.allreduce_episode_stats(env.stats, ctx=env.ctx)
    assert torch.equal(got, env.stats) and got.data_ptr() != env.stats.data_ptr()
    assert got[0] > 0 and got[15] == 4096 * 30
