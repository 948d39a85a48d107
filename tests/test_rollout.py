"""GAE over a device-resident rollout (acb200_gae) against the numpy restatement of SB3's
RolloutBuffer.compute_returns_and_advantage (oracle/rollout_oracle.py).  Bar: bit-exact (float32, same operation order)."""
import numpy as np
import pytest
import torch

from oracle import rollout_oracle as ro


def _case(T, n, seed, p_start=0.05):
    r = np.random.default_rng(seed)
    rewards = r.normal(0.5, 1.5, (T, n)).astype(np.float32)
    values = r.normal(5.0, 10.0, (T, n)).astype(np.float32)
    starts = (r.random((T, n)) < p_start).astype(np.float32)
    last = r.normal(5.0, 10.0, n).astype(np.float32)
    dones = r.random(n) < p_start
    return rewards, values, starts, last, dones


def test_oracle_against_scalar_float64_recursion():
    """Pins the restatement: the textbook recursion per environment in float64 agrees to float32 rounding, and an
    episode boundary cuts the bootstrap exactly where SB3 cuts it (episode_starts[t+1], then dones)."""
    T, n, g, lam = 40, 16, 0.99, 0.95
    rewards, values, starts, last, dones = _case(T, n, 1, p_start=0.15)
    adv, ret = ro.compute_returns_and_advantage(rewards, values, starts, last, dones, g, lam)
    for e in range(n):
        gae = 0.0
        for t in reversed(range(T)):
            nnt = 1.0 - (float(dones[e]) if t == T - 1 else float(starts[t + 1, e]))
            nv = float(last[e]) if t == T - 1 else float(values[t + 1, e])
            delta = float(rewards[t, e]) + g * nv * nnt - float(values[t, e])
            gae = delta + g * lam * nnt * gae
            assert abs(adv[t, e] - gae) <= 2e-5 * max(1.0, abs(gae))
            assert abs(ret[t, e] - (gae + float(values[t, e]))) <= 2e-5 * max(1.0, abs(gae))
    assert adv.dtype == np.float32 and ret.dtype == np.float32


@pytest.mark.gpu
@pytest.mark.parametrize("T,n", [(1, 1), (7, 300), (128, 4096), (2048, 8)])
def test_gae_bit_exact_vs_oracle(T, n):
    import acb200
    rewards, values, starts, last, dones = _case(T, n, T * 7 + n)
    buf = acb200.RolloutBufferB200(T, n, gamma=0.99, gae_lambda=0.95)
    buf.rewards.copy_(torch.from_numpy(rewards)); buf.values.copy_(torch.from_numpy(values))
    buf.episode_starts.copy_(torch.from_numpy(starts.astype(np.uint8)))
    buf.compute_returns_and_advantage(torch.from_numpy(last).cuda(), torch.from_numpy(dones).cuda())
    adv, ret = ro.compute_returns_and_advantage(rewards, values, starts, last, dones, 0.99, 0.95)
    assert np.array_equal(buf.advantages.cpu().numpy().view(np.uint32), adv.view(np.uint32))
    assert np.array_equal(buf.returns.cpu().numpy().view(np.uint32), ret.view(np.uint32))


@pytest.mark.gpu
def test_rollout_buffer_with_env_and_policy():
    """add() from the device env + device policy, GAE, and get(): shapes, environment-major flattening, and the property
    that an environment's advantages do not depend on the others (full-size columns repeat a 256-column run)."""
    import acb200
    T, n = 16, 1024
    env = acb200.BatchedRateControlEnv(n, controller="external", precision=32, seed=5, episode_length=0.2)
    pol = acb200.LearnedPolicyB200(acb200.policy.random_policy_state_dict("lstm", 0), kind="lstm", n=n, ctx=env.ctx)
    buf = acb200.RolloutBufferB200(T, n, ctx=env.ctx)
    env.reset()
    starts = torch.ones(n, dtype=torch.uint8, device="cuda")
    for t in range(T):
        obs = env.obs.clone()
        act = pol(obs, starts).clone()
        value = obs[:, 10] * 0.01                                  # any critic: here a function of altitude
        env.step(act)
        buf.add(obs, act, env.reward, starts, value)
        starts = (env.terminated | env.truncated).clone()
    assert buf.full
    buf.compute_returns_and_advantage(env.obs[:, 10] * 0.01, starts)
    adv, ret = ro.compute_returns_and_advantage(buf.rewards.cpu().numpy(), buf.values.cpu().numpy(),
                                                buf.episode_starts.float().cpu().numpy(), (env.obs[:, 10] * 0.01).cpu().numpy(),
                                                starts.bool().cpu().numpy())
    assert np.array_equal(buf.advantages.cpu().numpy(), adv) and np.array_equal(buf.returns.cpu().numpy(), ret)
    assert float(buf.episode_starts.float().mean()) > 0.05          # 0.2 s episodes: boundaries inside the rollout
    batches = list(buf.get(batch_size=4096))
    assert len(batches) == T * n // 4096 and batches[0]["observations"].shape == (4096, 18)
    assert batches[0]["observations"].is_cuda
    with pytest.raises(RuntimeError):
        buf.add(env.obs, act, env.reward, starts, value)
