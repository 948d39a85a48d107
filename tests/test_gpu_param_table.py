"""Per-aircraft AircraftParams (second half): acb200_build_param_table / acb200_set_param_table and the
table variant of the env step kernel, against ONE CPU oracle env per aircraft (its own params, id_offset = its id)."""
import ctypes as C

import numpy as np
import pytest
import torch

from conftest import rel_err
from oracle import oracle as orc

pytestmark = pytest.mark.gpu
TOL = 1e-9


def test_every_aircraft_matches_its_own_oracle_env():
    import acb200
    n, steps = 256, 120
    base = acb200.default_aircraft_params()
    params = acb200.envs.randomize_aircraft_params_table(base, n, spread=0.25, seed=3)
    assert len({bytes(p) for p in params}) == n                                    # all different
    env = acb200.BatchedRateControlEnv(n, controller="attitude_pid", precision=64, seed=11, episode_length=1.0, track_components=True)
    env.set_param_table(params)
    rng = np.random.default_rng(5)
    att = np.zeros((4, n)); att[0] = rng.uniform(-0.35, 0.35, n); att[1] = rng.uniform(-0.03, 0.08, n)
    att[2] = np.inf; att[3] = rng.uniform(0.5, 0.8, n)
    env.att_cmd.copy_(torch.as_tensor(att))
    refs = []
    for i in range(n):
        r = orc.Env(1, controller=acb200.envs.CONTROLLERS["attitude_pid"], seed=11, difficulty=0.5, command_type=0, episode_length=1.0,
                    params=orc.AircraftParams.from_buffer_copy(bytes(params[i])), id_offset=i, nthreads=1)
        r.att_cmd[:, 0] = att[:, i]
        r.reset()
        refs.append(r)
    env.reset()
    ok = np.ones(n, bool)
    resets = diverged = 0
    for t in range(steps):
        obs, rew, term, trunc = env.step()
        gs, gp = env.state.cpu().numpy(), env.pid.cpu().numpy().view(np.uint32)
        gt, gu, gr, go = term.cpu().numpy(), trunc.cpu().numpy(), rew.cpu().numpy(), obs.cpu().numpy()
        for i, r in enumerate(refs):
            robs, rrew, rterm, rtrunc = r.step()
            same = bool(np.all(gp[:, i] == r.pid.view(np.uint32)[:, 0]) and gt[i] == rterm[0] and gu[i] == rtrunc[0])
            if ok[i] and not same:
                diverged += 1
            ok[i] &= same
            if ok[i]:
                assert rel_err(gs[:, i], r.state[:, 0]) <= TOL, f"step {t + 1} aircraft {i}"
                assert abs(gr[i] - rrew[0]) <= 1e-7 and rel_err(go[i], robs[0]) <= 2e-7
            done = bool(rterm[0] or rtrunc[0])
            resets += done
            if done and gt[i] == rterm[0] and gu[i] == rtrunc[0]:
                ok[i] = True
    assert diverged <= n // 50 and ok.mean() > 0.98 and resets >= n
    # really different aircraft: a plain env with the same seeds and commands flies elsewhere
    plain = acb200.BatchedRateControlEnv(n, controller="attitude_pid", precision=64, seed=11, episode_length=1.0)
    plain.att_cmd.copy_(torch.as_tensor(att)); plain.reset()
    env.reset()                                     # episode 2 of `env` vs episode 1 of `plain` would differ anyway: compare dynamics
    env.state.copy_(plain.state); env.derived.copy_(plain.derived); env.pid.zero_(); plain.pid.zero_()
    for _ in range(10):
        env.step(); plain.step()
    assert not torch.allclose(env.state[:12], plain.state[:12], rtol=0, atol=1e-6)


@pytest.mark.parametrize("precision", [32, 64])
def test_table_of_identical_rows_equals_the_constant_bank_kernel(precision):
    """Same folded values, same arithmetic: bit-identical to the default path (the fp32 default is the pair kernel, which the
    one-aircraft kernels equal bit for bit), through step, step_host and a ranged step, on a ragged fleet."""
    import acb200
    n = 1000
    cessna = acb200.default_aircraft_params("cessna")
    a = acb200.BatchedRateControlEnv(n, controller="rate_pid", precision=precision, seed=2)
    b = acb200.BatchedRateControlEnv(n, controller="rate_pid", precision=precision, seed=2)
    a.ctx.check(a.ctx.lib.acb200_set_aircraft_params(a.ctx.handle, C.byref(cessna)))
    b.set_param_table([cessna] * n)
    a.reset(); b.reset()
    for _ in range(25):
        a.step(); b.step()
    for name in ("state", "obs", "reward", "pid", "rew_mem", "terminated", "truncated"):
        assert torch.equal(getattr(a, name), getattr(b, name)), name
    pin = lambda *shape, dt=torch.float32: torch.empty(shape, dtype=dt).pin_memory()
    obs_h, rew_h, te_h, tr_h = pin(n, 18), pin(n), pin(n, dt=torch.uint8), pin(n, dt=torch.uint8)
    a.step(); b.step_host(None, obs_h, rew_h, te_h, tr_h)
    assert torch.equal(a.state, b.state) and torch.equal(a.obs.cpu(), obs_h)
    lib = b.ctx.lib
    a.step()
    b.ctx.check(lib.acb200_env_step_range(b.ctx.handle, C.byref(b.c), C.byref(b.cfg), None, 0, 512))
    b.ctx.check(lib.acb200_env_step_range(b.ctx.handle, C.byref(b.c), C.byref(b.cfg), None, 512, n - 512))
    assert torch.equal(a.state, b.state)
    b.set_param_table(None)                                          # back to the context's own (default) parameters
    c = acb200.BatchedRateControlEnv(n, controller="rate_pid", precision=precision, seed=2)
    c.load_state_dict(b.state_dict())
    b.step(); c.step()
    assert torch.equal(b.state, c.state)


def test_param_table_argument_errors():
    import acb200
    n = 256
    env = acb200.BatchedRateControlEnv(n, controller="rate_pid", precision=32)
    base = acb200.default_aircraft_params()
    with pytest.raises(ValueError, match="one AircraftParams per aircraft"):
        env.set_param_table([base] * 10)
    bad = acb200.default_aircraft_params(); bad.inertia_yy = 0.0
    with pytest.raises(ValueError, match=r"inertia.*row 7"):
        env.set_param_table([base] * 7 + [bad] + [base] * (n - 8))
    steep = acb200.default_aircraft_params(); steep.max_alpha = 60.0
    with pytest.raises(ValueError, match="max_alpha"):               # the fp32 build's range, checked per row at build time
        env.set_param_table([steep] * n)
    env.reset(); env.step()                                          # nothing was set by the failed calls
    assert getattr(env, "param_table", None) is None
    env.set_param_table([base] * n)
    with pytest.raises(ValueError, match="groups"):
        env.set_param_groups([base, base])
    other = acb200.BatchedRateControlEnv(512, controller="rate_pid", precision=32, ctx=env.ctx)
    other.reset()
    with pytest.raises(ValueError, match="another fleet size"):
        other.step()
    assert env.param_table.shape == (env.ctx.lib.acb200_param_table_words(), n) and env.param_table.shape[0] == 79
