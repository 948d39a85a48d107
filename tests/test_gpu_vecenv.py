"""SB3 VecEnv protocol of B200VecEnv, driven by a stand-in for SB3's rollout loop (SB3 itself is not installed):
auto-reset semantics, terminal_observation, TimeLimit.truncated, Monitor-style episode dicts, curriculum switch."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_vecenv_protocol_and_auto_reset_semantics():
    import acb200
    from oracle import oracle as orc
    n = 512
    venv = acb200.B200VecEnv(n, difficulty="hard", episode_length=0.5, precision=64, seed=5, reward_components=True)   # 25-step episodes
    ref = orc.Env(n, controller=orc.CTRL_EXTERNAL, seed=5, difficulty=0.7, episode_length=0.5)
    assert venv.observation_space.shape == (18,) and venv.action_space.shape == (4,)
    assert venv.observation_space.low[9] == 5.0 and venv.action_space.low[3] == 0.0
    obs = venv.reset(); robs = ref.reset()
    assert obs.shape == (n, 18) and obs.dtype == np.float32 and np.array_equal(obs, robs)
    rng = np.random.default_rng(0)
    ret, length = np.zeros(n), np.zeros(n, int)
    episodes = 0
    for t in range(60):
        a = rng.uniform(-1, 1, (n, 4)).astype(np.float32) * np.array([0.3, 0.3, 0.3, 1.0], np.float32)
        venv.step_async(a)
        obs, rew, dones, infos = venv.step_wait()
        robs, rrew, rterm, rtrunc = ref.step(a)
        assert obs.dtype == np.float32 and rew.dtype == np.float32 and dones.dtype == bool and len(infos) == n
        assert np.array_equal(dones, (rterm | rtrunc).astype(bool))
        assert np.allclose(obs, robs, rtol=0, atol=2e-6) and np.allclose(rew, rrew, rtol=0, atol=1e-4)
        ret += rew; length += 1
        for i in np.flatnonzero(dones):
            info = infos[i]
            assert np.allclose(info["terminal_observation"], ref.terminal_obs[i], atol=2e-6)
            assert info["TimeLimit.truncated"] == bool(rtrunc[i] and not rterm[i])
            assert info["episode"]["l"] == length[i] and abs(info["episode"]["r"] - ret[i]) < 1e-3
            assert set(info["reward_components"]) >= {"tracking", "smoothness", "stability", "oscillation", "survival"}
            ret[i] = 0; length[i] = 0; episodes += 1
        for i in np.flatnonzero(~dones)[:3]:
            assert "terminal_observation" not in infos[i] and "episode" not in infos[i]
    assert episodes >= 2 * n
    assert venv.episode_stats()["episodes"] == episodes
    # curriculum switch without re-creating the env (train_rate.py:150-180 re-creates)
    venv.set_attr("difficulty", "easy")
    assert set(venv.get_attr("difficulty_scale", [0, 1, 2])) == {float(np.float32(0.3))}
    venv.set_attr("command_type", "sine", indices=[0, 1])
    assert venv.env.command_type[:3].tolist() == [2, 2, 0]
    with pytest.raises(ValueError):
        venv.step_async(np.zeros((n, 3), np.float32))
    venv.close()


def test_vecenv_lazy_infos_and_internal_controller():
    import acb200
    n = 8192
    venv = acb200.B200VecEnv(n, controller="rate_pid", episode_length=0.2, seed=1)
    venv.reset()
    for _ in range(10):
        obs, rew, dones, infos = venv.step(None)
    assert isinstance(infos, acb200.vecenv.LazyInfos) and len(infos) == n
    assert dones.all()                                   # 10-step episodes: everyone truncates (or crashed) on step 10
    i = int(np.flatnonzero(dones)[0])
    assert infos[i]["episode"]["l"] in range(1, 11) and infos[i]["terminal_observation"].shape == (18,)


def test_vecenv_infos_carry_the_reference_info_keys():
    """rate_env.py:410-435: time, step, position, rate_command, rate_error, airspeed, altitude, is_settled on every info."""
    import acb200
    n = 300
    venv = acb200.B200VecEnv(n, controller="rate_pid", precision=64, seed=4, episode_length=0.2, reward_components=True)   # 10-step episodes
    venv.reset()
    keys = {"time", "step", "position", "rate_command", "rate_error", "airspeed", "altitude", "is_settled", "reward_components"}
    for t in range(12):
        obs, rew, dones, infos = venv.step(None)
        assert len(infos) == n
        for i in (0, 7, n - 1):
            info = infos[i]
            assert keys <= set(info)
            if dones[i]:
                assert info["step"] == info["episode"]["l"] and np.isnan(info["position"]).all()
                assert np.allclose(info["rate_command"], info["terminal_observation"][3:6])
            else:
                assert info["step"] == (t % 10) + 1 and abs(info["time"] - info["step"] * 0.02) < 1e-12
                assert np.allclose(info["rate_command"], obs[i, 3:6], atol=1e-6) and np.allclose(info["rate_error"], obs[i, 6:9], atol=1e-6)
                assert abs(info["airspeed"] - obs[i, 9]) < 1e-4 and abs(info["altitude"] - obs[i, 10]) < 1e-3
    assert dones.any() or t > 0
