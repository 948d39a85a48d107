"""acb200_env_gather_done_noalloc: the allocation-free variant of acb200_env_gather_done (static device scratch, no
counterpart in the reference).  Called through the C ABI (ctypes) and compared
with the existing entry point -- ids are integers and the observations are copies, so the bar is bit-exact, as in
tests/test_gpu_env_tools.py -- on two calls in a row (the static array keeps the first call's contents), then on a smaller
fleet (stale counts of the larger one must not matter).  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _gather(env, fn, cap):
    d = env.device
    ids = torch.full((cap,), -1, device=d, dtype=torch.int64)
    tobs = torch.full((cap, 18), -7.0, device=d, dtype=torch.float32)
    robs = torch.full((cap, 18), -7.0, device=d, dtype=torch.float32)
    count = torch.full((1,), -1, device=d, dtype=torch.int64)
    env.ctx.set_stream(torch.cuda.current_stream(d).cuda_stream)
    rc = fn(env.ctx.handle, C.byref(env.c), cap, C.c_void_p(ids.data_ptr()), C.c_void_p(tobs.data_ptr()),
            C.c_void_p(robs.data_ptr()), C.c_void_p(count.data_ptr()))
    assert rc == 0, env.ctx.lib.acb200_last_error(env.ctx.handle)
    torch.cuda.synchronize()
    return int(count.item()), ids.cpu(), tobs.cpu(), robs.cpu()


def _make(n, seed):
    import acb200
    env = acb200.BatchedRateControlEnv(n, controller="rate_pid", precision=32, seed=seed, episode_length=0.4)    # 20-step episodes
    env.reset()
    env.step_count.copy_(torch.as_tensor((np.arange(n) * 7) % 20, dtype=torch.int32))                            # staggered truncations
    return env


def test_noalloc_gather_matches_the_allocating_entry_point_on_consecutive_calls():
    big, small = _make(70_001, seed=8), _make(3_000, seed=9)
    lib = big.ctx.lib
    assert lib.acb200_noalloc_max_aircraft() == 16384 * 1024
    calls = 0
    for env, steps in ((big, 2), (small, 2), (big, 1)):        # two calls in a row per fleet; small after big; big again
        for _ in range(steps):
            obs, rew, term, trunc = env.step()
            cap = env.n
            want = _gather(env, lib.acb200_env_gather_done, cap)
            free_before = torch.cuda.mem_get_info(env.device)[0]
            got = _gather(env, lib.acb200_env_gather_done_noalloc, cap)
            if calls > 0:                                      # (the very first call may load the kernels' module)
                assert torch.cuda.mem_get_info(env.device)[0] >= free_before, "the no-alloc entry point took device memory"
            calls += 1
            k = want[0]
            exp = torch.nonzero((term | trunc).bool().cpu()).flatten()
            assert got[0] == k == len(exp) and k > 0
            assert torch.equal(got[1][:k], want[1][:k]) and torch.equal(got[1][:k], exp)
            assert torch.equal(got[2][:k], want[2][:k]) and torch.equal(got[3][:k], want[3][:k])
            assert torch.equal(got[1][k:], want[1][k:]) and (got[1][k:] == -1).all()                # nothing written beyond count
    # a capacity below the count: the count is still exact, only `capacity` rows are written
    small.step_count.fill_(19); small.step()
    count, ids, tobs, _ = _gather(small, lib.acb200_env_gather_done_noalloc, 5)
    assert count == small.n and ids.tolist() == [0, 1, 2, 3, 4]


def test_noalloc_gather_rejects_what_it_cannot_hold():
    import acb200
    env = _make(256, seed=1)
    lib, h = env.ctx.lib, env.ctx.handle
    too_big = acb200._cabi.EnvC.from_buffer_copy(bytes(env.c))
    too_big.fleet.n = lib.acb200_noalloc_max_aircraft() + 1
    out = torch.zeros((4,), device="cuda", dtype=torch.int64)
    p = C.c_void_p(out.data_ptr())
    assert lib.acb200_env_gather_done_noalloc(h, C.byref(too_big), 4, p, None, None, p) == acb200._cabi.EVALUE
    msg = lib.acb200_last_error(h).decode()
    assert "16777216" in msg and "acb200_env_gather_done" in msg
    assert lib.acb200_env_gather_done_noalloc(h, C.byref(env.c), -1, p, None, None, p) == acb200._cabi.EVALUE
    assert lib.acb200_env_gather_done_noalloc(h, C.byref(env.c), 4, None, None, None, p) == acb200._cabi.ENULL
    # an empty fleet: count 0, no launch
    empty = acb200._cabi.EnvC.from_buffer_copy(bytes(env.c)); empty.fleet.n = 0
    before = env.ctx.launch_count
    out.fill_(9)
    assert lib.acb200_env_gather_done_noalloc(h, C.byref(empty), 4, p, None, None, p) == 0
    torch.cuda.synchronize()
    assert int(out[0].item()) == 0 and env.ctx.launch_count == before
