"""acb200_env_gather_done (compact list of finished aircraft) and the env checkpoint (state_dict / load_state_dict)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n", [1, 1000, 5000, 70_001])
def test_gather_done_lists_the_finished_aircraft_in_id_order(n):
    import acb200
    env = acb200.BatchedRateControlEnv(n, controller="rate_pid", precision=32, seed=8, episode_length=0.4)   # 20-step episodes
    env.set_curriculum(torch.as_tensor(np.array([0.3, 0.5, 0.7], np.float32)[np.arange(n) % 3]), "step")
    env.reset()
    env.step_count.copy_(torch.as_tensor((np.arange(n) * 7) % 20, dtype=torch.int32))      # stagger the truncations
    total = 0
    for t in range(25):
        obs, rew, term, trunc = env.step()
        ids, tobs, robs = env.gather_done()
        done = (term | trunc).bool().cpu()
        exp = torch.nonzero(done).flatten()
        assert torch.equal(ids, exp), t
        assert torch.equal(tobs, env.terminal_obs.cpu()[exp]) and torch.equal(robs, obs.cpu()[exp])
        total += len(ids)
    assert total >= n
    # capacity smaller than the number finished: the count is still right, the wrapper grows and retries
    env.step_count.fill_(19); env.step()
    ids, tobs, robs = env.gather_done(capacity=1)
    assert len(ids) == n and torch.equal(ids, torch.arange(n))


def test_gather_done_with_nobody_finished():
    import acb200
    env = acb200.BatchedRateControlEnv(300, controller="attitude_pid", precision=64, seed=1)
    env.att_cmd[3] = 0.6
    env.reset(); env.step()
    ids, tobs, robs = env.gather_done()
    assert len(ids) == 0 and tobs.shape == (0, 18)


@pytest.mark.parametrize("precision", [64, 32])
def test_checkpoint_resumes_bit_for_bit(precision):
    import acb200
    n = 2048
    mk = lambda seed: acb200.BatchedRateControlEnv(n, controller="rate_pid", precision=precision, seed=seed, episode_length=1.0)
    a = mk(5)
    a.set_curriculum(torch.as_tensor(np.array([0.3, 0.5, 0.7], np.float32)[np.arange(n) % 3]),
                     torch.as_tensor((np.arange(n) % 4).astype(np.uint8)))
    a.reset()
    for _ in range(70):
        a.step()
    sd = a.state_dict()
    assert all(not v.is_cuda for k, v in sd.items() if k != "config")
    b = mk(999)                                   # different seed: the checkpoint carries the Philox key and counters
    b.load_state_dict(sd)
    for _ in range(80):                           # crosses auto-resets (50-step episodes) and random-walk draws
        a.step(); b.step()
    for k in a._STATE_TENSORS:
        if k == "stats":        # sums of doubles built with cross-block atomics: the order, hence the last bits, varies per launch
            assert torch.allclose(a.stats, b.stats, rtol=1e-12, atol=1e-9), k
        else:
            assert torch.equal(getattr(a, k), getattr(b, k)), k
    with pytest.raises(ValueError):
        acb200.BatchedRateControlEnv(n + 1, precision=precision).load_state_dict(sd)


def test_dlpack_intake():
    """Anything with __dlpack__ is accepted where a tensor is taken (zero-copy for CUDA producers)."""
    import acb200

    class Foreign:                                # stands in for a CuPy / JAX array: only the DLPack protocol
        def __init__(self, t): self._t = t
        def __dlpack__(self, stream=None): return self._t.__dlpack__()
        def __dlpack_device__(self): return self._t.__dlpack_device__()

    n = 256
    fleet = acb200.AircraftFleet(n, precision=64)
    twin = acb200.AircraftFleet(n, precision=64)
    cmd = torch.zeros((4, n), device="cuda", dtype=torch.float64); cmd[0] = 0.3; cmd[3] = 0.7
    fleet.step_rate(Foreign(cmd), 0.01); twin.step_rate(cmd, 0.01)
    assert torch.equal(fleet.state, twin.state)
    env = acb200.BatchedRateControlEnv(n, precision=32, seed=1); env2 = acb200.BatchedRateControlEnv(n, precision=32, seed=1)
    env.reset(); env2.reset()
    act = torch.rand((n, 4), device="cuda") * 0.1
    env.step(Foreign(act)); env2.step(act)
    assert torch.equal(env.obs, env2.obs)
    env.step(act.cpu().numpy()); env2.step(act)   # host arrays are copied in
    assert torch.equal(env.obs, env2.obs)
