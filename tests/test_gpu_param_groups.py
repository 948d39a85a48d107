"""Per-group parameter randomisation: acb200_set_aircraft_params + acb200_env_step_range per
group of aircraft, against one CPU oracle env per group (its own AircraftParams, id_offset = the group's first id)."""
import ctypes as C

import numpy as np
import pytest
import torch

from conftest import rel_err
from helpers import rel_err_cols
from oracle import oracle as orc

pytestmark = pytest.mark.gpu
TOL = 1e-9


def _oracle_params(p):
    return orc.AircraftParams.from_buffer_copy(bytes(p))        # same 39 doubles, same order (tests/test_cabi_host.py)


def test_groups_match_one_oracle_env_per_group():
    import acb200
    n, bounds, steps = 512, [0, 128, 256, 384, 512], 150
    base = acb200.default_aircraft_params()
    groups = acb200.envs.randomize_aircraft_params(base, 4, spread=0.25, seed=3)
    assert bytes(groups[0]) == bytes(base) and bytes(groups[1]) != bytes(base)
    env = acb200.BatchedRateControlEnv(n, controller="attitude_pid", precision=64, seed=11, episode_length=1.0)
    env.set_param_groups(groups, bounds)
    rng = np.random.default_rng(5)
    att = np.zeros((4, n)); att[0] = rng.uniform(-0.35, 0.35, n); att[1] = rng.uniform(-0.03, 0.08, n)
    att[2] = np.inf; att[3] = rng.uniform(0.5, 0.8, n)
    env.att_cmd.copy_(torch.as_tensor(att))
    refs = []
    for g, (lo, hi) in enumerate(zip(bounds, bounds[1:])):
        r = orc.Env(hi - lo, controller=acb200.envs.CONTROLLERS["attitude_pid"], seed=11, difficulty=0.5, command_type=0,
                    episode_length=1.0, params=_oracle_params(groups[g]), id_offset=lo)
        r.att_cmd[:] = att[:, lo:hi]
        r.reset()
        refs.append(r)
    env.reset()
    resets = diverged = 0
    ok = np.ones(n, bool)
    for t in range(steps):
        obs, rew, term, trunc = env.step()
        for r, (lo, hi) in zip(refs, zip(bounds, bounds[1:])):
            robs, rrew, rterm, rtrunc = r.step()
            gterm, gtrunc = term[lo:hi].cpu().numpy(), trunc[lo:hi].cpu().numpy()
            # an aircraft stays in the comparison while its fp32 PID memories are bit-identical to the oracle's (as in
            # tests/test_gpu_env_f64.py); it rejoins at its next reset from the shared Philox stream
            same = (np.all(env.pid[:, lo:hi].cpu().numpy().view(np.uint32) == r.pid.view(np.uint32), axis=0)
                    & (gterm == rterm) & (gtrunc == rtrunc))
            k = ok[lo:hi]
            diverged += int((k & ~same).sum())
            k &= same
            e = rel_err_cols(env.state[:, lo:hi].cpu().numpy(), r.state)
            assert (e[k] <= TOL).all(), f"step {t + 1} group {lo}: {e[k].max():.3e}"
            assert np.max(np.abs(rew[lo:hi].cpu().numpy()[k] - rrew[k])) <= 1e-7
            assert rel_err(obs[lo:hi].cpu().numpy()[k], robs[k]) <= 2e-7
            done = (rterm | rtrunc).astype(bool)
            resets += int(done.sum())
            k |= done & (gterm == rterm) & (gtrunc == rtrunc)
            ok[lo:hi] = k
    assert diverged <= n // 50 and ok.mean() > 0.98
    assert resets >= n                                              # every aircraft went through the in-kernel auto-reset
    # the groups really fly different aircraft: same commands, same seeds, different trajectories
    plain = acb200.BatchedRateControlEnv(n, controller="attitude_pid", precision=64, seed=11, episode_length=1.0)
    plain.att_cmd.copy_(torch.as_tensor(att)); plain.reset()
    for t in range(20):
        plain.step()
    env2 = acb200.BatchedRateControlEnv(n, controller="attitude_pid", precision=64, seed=11, episode_length=1.0)
    env2.set_param_groups(groups, bounds); env2.att_cmd.copy_(torch.as_tensor(att)); env2.reset()
    for t in range(20):
        env2.step()
    assert torch.equal(env2.state[:, :128], plain.state[:, :128])                   # group 0 is the base aircraft
    assert not torch.allclose(env2.state[:12, 128:], plain.state[:12, 128:], rtol=0, atol=1e-6)


def test_one_group_equals_the_plain_step_and_fp32_runs():
    import acb200
    n = 1000                                                         # ragged tail in the last group
    base = acb200.default_aircraft_params()
    a = acb200.BatchedRateControlEnv(n, controller="rate_pid", precision=32, seed=2)
    b = acb200.BatchedRateControlEnv(n, controller="rate_pid", precision=32, seed=2)
    b.set_param_groups([base, base, base])                           # default even split: 384, 384, 232
    a.reset(); b.reset()
    l0 = b.ctx.launch_count
    for _ in range(30):
        a.step(); b.step()
    assert b.ctx.launch_count - l0 == 90                             # one launch per group and step
    assert torch.equal(a.state, b.state) and torch.equal(a.obs, b.obs) and torch.equal(a.reward, b.reward)
    b.set_param_groups(None)
    a.step(); b.step()
    assert torch.equal(a.state, b.state)


def test_argument_errors():
    import acb200
    env = acb200.BatchedRateControlEnv(512, controller="rate_pid", precision=32)
    base = acb200.default_aircraft_params()
    with pytest.raises(ValueError):
        env.set_param_groups([base, base], [0, 100, 512])            # not a multiple of 128
    with pytest.raises(ValueError):
        env.set_param_groups([base, base], [0, 256, 500])            # does not end at n
    bad = acb200.default_aircraft_params(); bad.mass = -1.0
    with pytest.raises(ValueError):
        env.set_param_groups([base, bad])                            # AircraftParams.__post_init__: "Mass must be positive"
    lib, h = env.ctx.lib, env.ctx.handle
    assert lib.acb200_env_step_range(h, C.byref(env.c), C.byref(env.cfg), None, 64, 64) == acb200._cabi.EVALUE
    assert lib.acb200_env_step_range(h, C.byref(env.c), C.byref(env.cfg), None, 384, 256) == acb200._cabi.EVALUE
    assert lib.acb200_env_step_range(h, C.byref(env.c), C.byref(env.cfg), None, 128, 0) == 0


def test_host_step_and_vecenv_honour_the_groups_and_failed_validation_changes_nothing():
    """the groups live in the context (acb200_set_param_groups), so step_host -- and every
    B200VecEnv.step -- flies the same aircraft as step(); a rejected group list leaves groups and parameters untouched."""
    import acb200
    n, bounds = 1024, [0, 256, 640, 1024]
    base = acb200.default_aircraft_params()
    groups = acb200.envs.randomize_aircraft_params(base, 3, spread=0.25, seed=9)
    mk = lambda: acb200.BatchedRateControlEnv(n, controller="rate_pid", precision=32, seed=4)
    a, b, plain = mk(), mk(), mk()
    a.set_param_groups(groups, bounds); b.set_param_groups(groups, bounds)
    bad = acb200.default_aircraft_params(); bad.mass = -1.0
    with pytest.raises(ValueError, match="Mass must be positive"):
        b.set_param_groups([groups[1], bad])                         # rejected as a whole
    got = acb200._cabi.AircraftParamsC()
    assert b.ctx.lib.acb200_get_aircraft_params(b.ctx.handle, C.byref(got)) == 0 and bytes(got) == bytes(base)
    a.reset(); b.reset(); plain.reset()
    pin = lambda *shape, dt=torch.float32: torch.empty(shape, dtype=dt).pin_memory()
    obs_h, rew_h, te_h, tr_h = pin(n, 18), pin(n), pin(n, dt=torch.uint8), pin(n, dt=torch.uint8)
    for _ in range(25):
        a.step(); plain.step()
        b.step_host(None, obs_h, rew_h, te_h, tr_h)
    torch.cuda.synchronize()
    assert torch.equal(a.state, b.state) and torch.equal(a.obs.cpu(), obs_h) and torch.equal(a.reward.cpu(), rew_h)
    assert torch.equal(a.state[:, :256], plain.state[:, :256]) and not torch.equal(a.state[:, 256:], plain.state[:, 256:])
    # outputs are optional: nothing but reward + flags leaves the device, results unchanged
    a.step(); b.step_host(None, None, rew_h, te_h, None)
    assert torch.equal(a.reward.cpu(), rew_h) and torch.equal(a.terminated.cpu(), te_h) and torch.equal(a.state, b.state)
    # a fleet of another size under the same context is refused, not silently simulated with the wrong split
    other = acb200.BatchedRateControlEnv(512, controller="rate_pid", precision=32, ctx=a.ctx)
    other.reset()
    with pytest.raises(ValueError):
        other.step()
