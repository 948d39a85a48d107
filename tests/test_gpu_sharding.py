"""Sharding invariance on one GPU: two half-fleets with id offsets give, bit for bit, what the whole fleet gives."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("precision", [64, 32])
def test_results_do_not_depend_on_the_sharding(precision):
    import torch
    import acb200
    from acb200 import sharding
    n, world = 6001, 2
    diff = np.array([0.3, 0.5, 0.7], np.float32)[np.arange(n) % 3]
    kw = dict(controller="rate_pid", precision=precision, seed=77, episode_length=0.3)
    whole = acb200.BatchedRateControlEnv(n, **kw)
    whole.set_curriculum(torch.as_tensor(diff), "step"); whole.reset()
    parts = []
    for r in range(world):
        first, count = sharding.shard_range(n, r, world)
        e = sharding.make_sharded_env(n, r, world, device=0, **kw)
        e.set_curriculum(torch.as_tensor(diff[first:first + count]), "step"); e.reset()
        parts.append((first, count, e))
    for t in range(40):        # 15-step episodes: two generations of in-kernel resets
        whole.step()
        for _, _, e in parts:
            e.step()
    for first, count, e in parts:
        assert torch.equal(e.state, whole.state[:, first:first + count])
        assert torch.equal(e.obs, whole.obs[first:first + count])
        assert torch.equal(e.reward, whole.reward[first:first + count])
        assert torch.equal(e.episode_idx, whole.episode_idx[first:first + count])
    tot = sum(e.stats for _, _, e in parts)
    assert tot[0] == whole.stats[0] and tot[15] == whole.stats[15] == 40 * n
    # the per-warp partial sums are formed in the arithmetic type of the build, so their grouping shows at float precision
    assert torch.allclose(tot, whole.stats, rtol=1e-9 if precision == 64 else 2e-6, atol=1e-6 if precision == 64 else 1e-2)
