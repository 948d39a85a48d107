"""RateControlEnvB200 / ResidualRateControlEnvB200: the reference's N = 1 gym.Env API, replayed on the
golden episodes of the unmodified reference env (tests/golden/rate_env.npz, rate_env_random.npz)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, rel_err

pytestmark = pytest.mark.gpu
INFO_KEYS = {"time", "step", "position", "rate_command", "rate_error", "airspeed", "altitude", "is_settled", "reward_components"}
DIFF = {0.3: "easy", 0.5: "medium", 0.7: "hard"}
CTYPE = {0: "step", 1: "ramp", 2: "sine", 3: "random"}


def _inject(env, g, k, orc, torch):
    """the reference's own initial draws (its RNG streams cannot be replayed)"""
    b = env.batched
    s0 = g[f"c{k}_s0"]
    b.state[:, 0] = torch.as_tensor(s0, device="cuda")
    b.derived[:, 0] = torch.as_tensor(orc.derived(np.ascontiguousarray(s0[:12])), device="cuda")
    if f"c{k}_cmd0" in g.files:
        env.rate_command = g[f"c{k}_cmd0"]
        b.sched[:, 0] = torch.as_tensor(g[f"c{k}_sched"], device="cuda")


@pytest.mark.parametrize("k", [0, 3, 4, 12, 15])
def test_gym_env_replays_reference_episodes(k):
    import torch
    import acb200
    from oracle import oracle as orc
    g = np.load(os.path.join(GOLDEN, "rate_env.npz"))
    seed, scale, ctype, ctrl = g[f"c{k}_meta"]
    cls = acb200.ResidualRateControlEnvB200 if int(ctrl) == 3 else acb200.RateControlEnvB200
    env = cls(difficulty=DIFF[float(scale)], command_type=CTYPE[int(ctype)], rng_seed=int(seed))
    obs, info = env.reset(seed=int(seed))
    assert obs.shape == (18,) and obs.dtype == np.float32 and set(info) >= INFO_KEYS - {"reward_components"}
    assert info["step"] == 0 and info["time"] == 0.0 and env.max_steps == 500 and env.dt == 0.02
    assert env.observation_space.shape == (18,) and env.action_space.shape == (4,)
    _inject(env, g, k, orc, torch)
    acts, T = g[f"c{k}_actions"], min(len(g[f"c{k}_reward"]), 120)
    t_acc = 0.0
    for t in range(T):
        obs, rew, term, trunc, info = env.step(acts[t])
        ctx = f"case {k} step {t + 1}"
        t_acc += 0.02
        assert isinstance(rew, float) and isinstance(term, bool) and isinstance(trunc, bool)
        assert set(info) == INFO_KEYS | ({"pid_action", "residual_action", "combined_action"} if int(ctrl) == 3 else set()), ctx
        assert rel_err(obs, g[f"c{k}_obs"][t]) <= 2e-7 and abs(rew - g[f"c{k}_reward"][t]) <= 1e-7, ctx
        assert term == bool(g[f"c{k}_terminated"][t]) and trunc == bool(g[f"c{k}_truncated"][t]), ctx
        st = g[f"c{k}_states"][t]
        assert info["step"] == t + 1 and info["time"] == t_acc, ctx                      # fp64 accumulation of dt, rate_env.py:241
        assert rel_err(info["position"], st[0:3]) <= 1e-9 and rel_err(info["rate_command"], g[f"c{k}_rate_cmd"][t]) <= 1e-12, ctx
        assert rel_err(info["rate_error"], g[f"c{k}_rate_cmd"][t] - st[9:12]) <= 1e-9, ctx
        assert info["is_settled"] == bool(g[f"c{k}_is_settled"][t]) == env.settle_bonus.is_settled, ctx
        rc, exp = info["reward_components"], g[f"c{k}_components"][t]
        for j, name in enumerate(("tracking", "smoothness", "stability", "oscillation", "survival")):
            assert abs(rc[name] - exp[j]) <= 1e-7, (ctx, name)
        assert abs(rc["total"] - exp[:5].sum()) <= 1e-7 and ("crash_penalty" in rc) == (term and not trunc)
        sim = env.sim.get_state()
        assert rel_err(sim.angular_rate, st[9:12]) <= 1e-9 and abs(sim.altitude - info["altitude"]) == 0.0
        if int(ctrl) == 3:
            assert np.allclose(info["combined_action"], env.prev_action) and info["residual_action"].dtype == np.float32
        if term or trunc:
            break
    assert abs(sum(env.episode_rewards[n] for n in ("tracking", "survival")) -
               g[f"c{k}_components"][:t + 1, [0, 4]].sum()) <= 1e-5


def test_gym_env_seeding_and_assignment():
    import acb200
    env = acb200.RateControlEnvB200(difficulty="hard", command_type="step", rng_seed=3)
    o1, _ = env.reset(seed=11); c1 = env.rate_command
    for _ in range(5):
        env.step(np.array([0.01, 0.0, 0.0, 0.6], np.float32))
    o2, _ = env.reset(seed=11)
    assert np.array_equal(o1, o2) and np.array_equal(c1, env.rate_command)           # equal seeds, equal episodes
    o3, _ = env.reset()
    assert not np.array_equal(o1, o3)                                                # the stream moves on without a seed
    env.rate_command = [0.5, 0.0, -0.2]                                              # eval_rate.py:104-108 assigns it
    obs, _, _, _, info = env.step(np.array([0, 0, 0, 0.6], np.float32))
    assert np.allclose(info["rate_command"], [0.5, 0.0, -0.2]) and np.allclose(obs[3:6], [0.5, 0.0, -0.2])
    with pytest.raises(KeyError):
        acb200.RateControlEnvB200(difficulty="impossible")
    with pytest.raises(ValueError):
        acb200.RateControlEnvB200(command_type="zigzag")
