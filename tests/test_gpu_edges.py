"""Edge cases of the C-ABI path: empty and ragged fleets, single aircraft, bad arguments, non-finite inputs."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("precision", [32, 64])
def test_ragged_sizes_match_a_larger_run(precision):
    """n = 1, 127, 129, 1000: every aircraft's result equals the same aircraft's result inside a 4096-aircraft run
    (block tails, warp tails, the obs staging through shared memory)."""
    import acb200
    big = acb200.BatchedRateControlEnv(4096, controller="attitude_pid", precision=precision, seed=9)
    big.reset()
    for _ in range(3):
        big.step()
    for n in (1, 127, 129, 1000):
        env = acb200.BatchedRateControlEnv(n, controller="attitude_pid", precision=precision, seed=9)
        env.reset()
        for _ in range(3):
            env.step()
        assert torch.equal(env.state, big.state[:, :n])
        assert torch.equal(env.obs, big.obs[:n])
        assert torch.equal(env.reward, big.reward[:n])


def test_empty_fleet_is_a_no_op():
    import acb200
    import ctypes as C
    ctx = acb200.Context(0)
    before = ctx.launch_count
    # GAE over nothing, policy over nothing: OK without a launch
    assert ctx.lib.acb200_gae(ctx.handle, 0, 0, 1, 1, 1, 1, 1, 0.99, 0.95, 1, 1) == 0
    pol = acb200.LearnedPolicyB200(acb200.policy.random_policy_state_dict("mlp", 0), kind="mlp", n=1, ctx=ctx)
    obs = torch.zeros((1, 18), device="cuda"); act = torch.zeros((1, 4), device="cuda")
    assert ctx.lib.acb200_policy_forward(ctx.handle, pol.handle, 0, obs.data_ptr(), None, None, None, None, act.data_ptr()) == 0
    assert ctx.launch_count == before


def test_bad_arguments_raise_like_the_reference():
    import acb200
    fleet = acb200.AircraftFleet(8, precision=64)
    ctrl = torch.zeros((4, 8), dtype=torch.float64, device="cuda")
    coarse = acb200.AircraftFleet(8, precision=64, config={"dt_physics": 2.0})
    with pytest.raises(ValueError):                      # simplified_6dof.py:241-245: substep above max_timestep
        coarse.step_surface(ctrl, 2.0)
    with pytest.raises(ValueError):
        fleet.step_surface(ctrl, float("nan"))
    with pytest.raises(ValueError):
        fleet.step_surface(ctrl, -0.01)
    with pytest.raises(RuntimeError):                    # mission not set
        fleet.step_waypoint(0.01)
    ctx = acb200.Context(0)
    assert ctx.lib.acb200_gae(ctx.handle, 4, 4, None, None, None, None, None, 0.99, 0.95, None, None) == -3   # ACB200_ENULL


def test_fp32_build_refuses_parameters_outside_its_polynomial_ranges():
    """The throughput build evaluates atan on |tan(alpha)| <= 1 (physics_f32.cuh): max_alpha beyond 45 degrees is an error
    there, not a silent loss of accuracy; the parity build takes what the reference takes."""
    import ctypes as C
    import acb200
    for precision, ok in ((32, False), (64, True)):
        env = acb200.BatchedRateControlEnv(64, controller="rate_pid", precision=precision, seed=1)
        env.reset()
        p = acb200.default_aircraft_params()
        p.max_alpha = 60.0
        env.ctx.check(env.ctx.lib.acb200_set_aircraft_params(env.ctx.handle, C.byref(p)))
        if ok:
            env.step()
        else:
            with pytest.raises(ValueError, match="max_alpha"):
                env.step()
        env.ctx.check(env.ctx.lib.acb200_set_aircraft_params(env.ctx.handle, C.byref(acb200.default_aircraft_params())))
        env.step()


def test_non_finite_observation_stays_in_its_row():
    """A NaN observation poisons that aircraft's action only (rows are independent through every layer)."""
    import acb200
    n = 256
    pol = acb200.LearnedPolicyB200(acb200.policy.random_policy_state_dict("lstm", 0), kind="lstm", n=n)
    obs = torch.randn(n, 18, device="cuda")
    clean = pol(obs).clone()
    pol.reset_state()
    obs2 = obs.clone(); obs2[5, 3] = float("nan"); obs2[200, 0] = float("inf")
    dirty = pol(obs2).clone()
    keep = torch.ones(n, dtype=torch.bool, device="cuda"); keep[5] = False; keep[200] = False
    assert torch.equal(dirty[keep], clean[keep])
    assert not torch.isfinite(dirty[5]).all()


def test_nan_action_freezes_the_poisoned_derivatives_in_both_builds(monkeypatch):
    """a NaN action passes np.clip and set_controls as NaN (rate_env.py:225, simplified_6dof.py:221-226); the
    reference's _dynamics then zeroes exactly the derivative words that control enters (nan_to_num, :496-501) -- those state
    words stay frozen over the control step, everything else integrates on.  The fp64 build does that op for op (checked
    against the oracle); the fp32 build takes a cold path that must agree with it to the fp32 per-step tolerance instead of
    saturating the accelerations (fminf/fmaxf clamps map NaN to a bound), counts the repairs, and stays bit-identical
    between the pair and the one-aircraft kernel."""
    import acb200
    from oracle import oracle as orc
    n = 512
    rng = np.random.default_rng(17)
    act = (rng.uniform(-1, 1, (n, 4)) * 0.05).astype(np.float32); act[:, 3] = 0.6
    slot = np.arange(n) % 8                       # slots 0..3: that control is NaN; 4: all four; 5..7: healthy aircraft
    for k in range(4):
        act[slot == k, k] = np.nan
    act[slot == 4] = np.nan
    a_t = torch.as_tensor(act, device="cuda")
    mk = lambda prec: acb200.BatchedRateControlEnv(n, controller="external", precision=prec, seed=21, auto_reset=False)
    e64, e32 = mk(64), mk(32)
    ref = orc.Env(n, controller=orc.CTRL_EXTERNAL, seed=21, auto_reset=False)
    e64.reset(); e32.reset(); ref.reset()
    before = e64.state.clone()
    for _ in range(2):
        e64.step(a_t); e32.step(a_t); ref.step(act)
    s64, s32 = e64.state.cpu().numpy(), e32.state.double().cpu().numpy()
    assert np.max(np.abs(s64 - ref.state) / np.maximum(np.abs(ref.state), 1.0)) <= 1e-9
    err = np.abs(s32 - s64) / np.maximum(np.abs(s64), 1.0)
    assert err.max() <= 2e-5, err.max()
    # frozen words: elevator (env action slot 1) -> u, w, q; aileron (slot 0) -> p; rudder (slot 2) -> v, r; throttle -> u
    b = before.cpu().numpy()
    for sl, words in ((0, [9]), (1, [3, 5, 10]), (2, [4, 11]), (3, [3]), (4, [3, 4, 5, 9, 10, 11])):
        m = slot == sl
        assert np.array_equal(s64[np.ix_(words, np.where(m)[0])], b[np.ix_(words, np.where(m)[0])])
        assert np.array_equal(e32.state.cpu().numpy()[np.ix_(words, np.where(m)[0])], before.float().cpu().numpy()[np.ix_(words, np.where(m)[0])])
    assert np.all(np.isfinite(s32))
    st = e32.episode_stats()
    assert st["nonfinite_repairs"] > 0
    # pair kernel == one-aircraft kernel, bit for bit, with NaN actions in one or both lanes
    monkeypatch.setenv("ACB200_ENV_PAIR", "0")
    e1 = mk(32); e1.reset()
    for _ in range(2):
        e1.step(a_t)
    for name in ("state", "obs", "prev_action", "derived"):
        x, y = getattr(e32, name), getattr(e1, name)
        assert torch.equal(torch.nan_to_num(x, nan=-77.0), torch.nan_to_num(y, nan=-77.0)), name
    rw, r1 = e32.reward, e1.reward
    assert torch.equal(torch.isnan(rw), torch.isnan(r1)) and torch.equal(rw[~torch.isnan(rw)], r1[~torch.isnan(r1)])
