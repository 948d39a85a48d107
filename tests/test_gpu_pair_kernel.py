"""The throughput build's default env step owns TWO aircraft per thread and runs the substeps on Blackwell's packed FP32
instructions (csrc/kernels_pair_f32.cuh, physics_f32x2.cuh).  Every lane operation is the IEEE operation of the
one-aircraft-per-thread kernel in the same order, so the two kernels must agree BIT FOR BIT on every buffer -- which also
puts the pair kernel under every fp32 tolerance the scalar kernel is tested against (tests/test_gpu_parity_f32.py)."""
import ctypes as C
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

BUFFERS = ("state", "controls", "derived", "pid", "rate_cmd", "prev_action", "rew_mem", "sched", "env_time", "step_count",
           "is_settled", "episode_idx", "ep_return", "obs", "reward", "terminated", "truncated", "terminal_obs", "components")


def _make(acb200, n, controller, seed, command_mix):
    env = acb200.BatchedRateControlEnv(n, controller=controller, precision=32, seed=seed, episode_length=1.0,
                                       track_components=True)
    g = torch.Generator().manual_seed(seed)
    if command_mix:
        env.set_curriculum(torch.tensor([0.3, 0.5, 0.7])[torch.randint(0, 3, (n,), generator=g)],
                           torch.randint(0, 4, (n,), generator=g).to(torch.uint8))
    att = torch.zeros(4, n)
    att[0].uniform_(-0.5, 0.5, generator=g); att[1].uniform_(-0.1, 0.15, generator=g)
    att[2] = float("inf"); att[3].uniform_(0.3, 0.9, generator=g)
    env.att_cmd.copy_(att)
    return env


def _run(acb200, pair, n, controller, steps, seed, command_mix, ranges=None):
    os.environ["ACB200_ENV_PAIR"] = "1" if pair else "0"
    try:
        env = _make(acb200, n, controller, seed, command_mix)
        env.reset()
        g = torch.Generator().manual_seed(seed + 1)
        rewards = []
        for t in range(steps):
            action = None
            if controller in ("external", "residual"):
                action = (torch.rand(n, 4, generator=g) * 2 - 1).mul_(0.6 if controller == "external" else 1.0).cuda()
                if controller == "external":
                    action[:, 3] = action[:, 3].abs()
            if ranges is None:
                env.step(action)
            else:
                env._bind_stream()
                for lo, cnt in ranges:
                    env.ctx.check(env.ctx.lib.acb200_env_step_range(env.ctx.handle, C.byref(env.c), C.byref(env.cfg),
                                                                    None if action is None else C.c_void_p(action.data_ptr()),
                                                                    lo, cnt))
            rewards.append(env.reward.clone())
        torch.cuda.synchronize()
        out = {k: getattr(env, k).cpu().numpy().copy() for k in BUFFERS}
        out["rewards"] = torch.stack(rewards).cpu().numpy()
        out["stats"] = env.stats.cpu().numpy().copy()
        return out
    finally:
        os.environ.pop("ACB200_ENV_PAIR", None)


def _same_bits(a, b, what):
    for k in a:
        if k == "stats":
            continue
        x, y = np.ascontiguousarray(a[k]), np.ascontiguousarray(b[k])
        assert x.shape == y.shape and x.dtype == y.dtype
        assert np.array_equal(x.view(np.uint8), y.view(np.uint8)), f"{what}: {k} differs in {(x != y).sum()} places"
    # episode statistics are float partial sums in a different association; the counters are exact
    sa, sb = a["stats"], b["stats"]
    assert np.array_equal(sa[[0, 3, 4, 5, 14, 15]], sb[[0, 3, 4, 5, 14, 15]])
    assert np.allclose(sa, sb, rtol=2e-5, atol=1e-3)


@pytest.mark.parametrize("controller", ["attitude_pid", "rate_pid", "external", "residual"])
def test_pair_kernel_is_bit_identical_to_the_one_aircraft_kernel(controller):
    import acb200
    n, steps = 2048 + 2 * 37, 130          # a ragged last block; 130 steps of a 50-step episode: every aircraft resets twice
    a = _run(acb200, False, n, controller, steps, seed=5, command_mix=True)
    b = _run(acb200, True, n, controller, steps, seed=5, command_mix=True)
    assert a["stats"][0] >= 2 * n                                  # the in-kernel auto-reset ran on both sides
    _same_bits(a, b, controller)


def test_pair_kernel_ranges_odd_counts():
    """env_step_range (first a multiple of 128) with odd counts (the pair kernel plus one scalar aircraft), a one- and a
    two-aircraft range, against the scalar kernel on the same ranges; whole-block ranges against one whole-fleet launch."""
    import acb200
    n = 1024
    ranges = [(0, 127), (128, 1), (256, 2), (384, 255), (640, 383)]
    a = _run(acb200, False, n, "attitude_pid", 60, seed=9, command_mix=False, ranges=ranges)
    b = _run(acb200, True, n, "attitude_pid", 60, seed=9, command_mix=False, ranges=ranges)
    _same_bits(a, b, "ranges")
    ranges = [(0, 128), (128, 384), (512, 512)]
    b = _run(acb200, True, n, "attitude_pid", 60, seed=9, command_mix=False, ranges=ranges)
    c = _run(acb200, True, n, "attitude_pid", 60, seed=9, command_mix=False)
    _same_bits(b, c, "ranges vs whole fleet")


def test_odd_fleet_size_falls_back_to_the_one_aircraft_kernel():
    import acb200
    a = _run(acb200, False, 777, "rate_pid", 40, seed=2, command_mix=True)
    b = _run(acb200, True, 777, "rate_pid", 40, seed=2, command_mix=True)
    _same_bits(a, b, "odd n")


def test_pair_kernel_repairs_nonfinite_state_like_the_one_aircraft_kernel():
    import acb200
    outs = []
    for pair in (False, True):
        os.environ["ACB200_ENV_PAIR"] = "1" if pair else "0"
        try:
            env = acb200.BatchedRateControlEnv(512, controller="rate_pid", precision=32, seed=1, auto_reset=False)
            env.reset()
            env.state[0, 3] = float("nan"); env.state[4, 10] = float("inf"); env.state[7, 11] = float("-inf")
            env.state[2, 200] = float("inf"); env.state[9, 201] = float("nan")
            for _ in range(3):
                env.step()
            torch.cuda.synchronize()
            outs.append((env.state.cpu().numpy().copy(), env.stats.cpu().numpy().copy()))
        finally:
            os.environ.pop("ACB200_ENV_PAIR", None)
    assert np.array_equal(outs[0][0].view(np.uint32), outs[1][0].view(np.uint32))
    assert np.isfinite(outs[1][0]).all() and outs[0][1][14] == outs[1][1][14] > 0


@pytest.mark.parametrize("level,dt", [("surface", 0.01), ("rate", 0.01), ("attitude", 0.02), ("surface", 0.001)])
def test_fleet_pair_kernels_are_bit_identical_to_the_one_aircraft_kernels(level, dt):
    """AircraftFleet.step_surface / step_rate / step_attitude (backend.set_controls + backend.step, with the agents in front)
    on aircraft pairs against the one-aircraft kernels: state, PID memory, held controls, derived outputs, surfaces."""
    import acb200
    n, steps = 1000, 120
    g = torch.Generator().manual_seed(4)
    init = torch.zeros(13, n)
    init[2].uniform_(-200, -50, generator=g); init[3].uniform_(15, 30, generator=g)
    init[6].uniform_(-0.3, 0.3, generator=g); init[7].uniform_(-0.2, 0.2, generator=g); init[8].uniform_(0, 6.28, generator=g)
    init[9:12].uniform_(-0.1, 0.1, generator=g)
    cmd = torch.zeros(4, n)
    if level == "surface":
        cmd[:3].uniform_(-0.3, 0.3, generator=g); cmd[3].uniform_(0.3, 0.9, generator=g)
    elif level == "rate":
        cmd[:3].uniform_(-0.5, 0.5, generator=g); cmd[3] = 0.7
    else:
        cmd[0].uniform_(-0.4, 0.4, generator=g); cmd[1].uniform_(-0.1, 0.15, generator=g); cmd[2] = float("inf"); cmd[3] = 0.6
    outs = []
    for pair in (False, True):
        os.environ["ACB200_ENV_PAIR"] = "1" if pair else "0"
        try:
            fleet = acb200.AircraftFleet(n, precision=32)
            fleet.reset(init.cuda())
            surf = torch.zeros(4, n, device="cuda")
            c = cmd.cuda()
            for _ in range(steps):
                if level == "surface":
                    fleet.step_surface(c, dt)
                elif level == "rate":
                    fleet.step_rate(c, dt, surfaces_out=surf)
                else:
                    fleet.step_attitude(c, dt, surfaces_out=surf)
            torch.cuda.synchronize()
            outs.append([t.cpu().numpy().copy() for t in (fleet.state, fleet.pid, fleet.controls, fleet.derived, surf)])
        finally:
            os.environ.pop("ACB200_ENV_PAIR", None)
    for x, y, name in zip(outs[0], outs[1], ("state", "pid", "controls", "derived", "surfaces")):
        assert np.array_equal(x.view(np.uint32), y.view(np.uint32)), f"{level}: {name} differs in {(x != y).sum()} places"
    assert np.isfinite(outs[1][0]).all() and np.ptp(outs[1][0][3]) > 1.0          # the fleet really flew


@pytest.mark.parametrize("level,dt,loop", [("hsa", 0.01, False), ("waypoint", 0.001, True), ("waypoint", 0.01, False)])
def test_guided_pair_kernels_are_bit_identical_to_the_one_aircraft_kernels(level, dt, loop):
    """Levels 2 and 1 (HSA / mission + waypoint guidance on top of the cascade) on aircraft pairs.  The non-looping mission
    is short enough that aircraft complete it at different steps: pairs with a finished aircraft take the cold path and
    the finished aircraft must stay untouched."""
    import acb200
    n, steps = 2000, 400 if dt == 0.01 else 300
    g = torch.Generator().manual_seed(8)
    init = torch.zeros(13, n)
    init[0].uniform_(-20, 20, generator=g); init[1].uniform_(-20, 20, generator=g); init[2] = -100.0
    init[3] = 15.0; init[8].uniform_(-0.2, 0.2, generator=g)
    cmd = torch.zeros(3, n)
    cmd[0].uniform_(-1.0, 1.0, generator=g); cmd[1].uniform_(13, 18, generator=g); cmd[2].uniform_(80, 120, generator=g)
    wps = [[25.0, 0.0, -100.0, 15.0], [50.0, 20.0, -100.0, 15.0]]
    outs = []
    for pair in (False, True):
        os.environ["ACB200_ENV_PAIR"] = "1" if pair else "0"
        try:
            fleet = acb200.AircraftFleet(n, precision=32)
            fleet.reset(init.cuda())
            surf = torch.zeros(4, n, device="cuda")
            hsa_cmd = torch.zeros(3, n, device="cuda")
            if level == "waypoint":
                fleet.set_mission(wps, acceptance_radius=12.0, guidance="PP", loop=loop)
            c = cmd.cuda()
            for _ in range(steps):
                if level == "hsa":
                    fleet.step_hsa(c, dt, surfaces_out=surf)
                else:
                    fleet.step_waypoint(dt, surfaces_out=surf, hsa_cmd_out=hsa_cmd)
            torch.cuda.synchronize()
            outs.append([t.cpu().numpy().copy() for t in (fleet.state, fleet.pid, fleet.pid_outer, fleet.controls, fleet.derived,
                                                           surf, hsa_cmd, fleet.wp_index)])
        finally:
            os.environ.pop("ACB200_ENV_PAIR", None)
    names = ("state", "pid", "pid_outer", "controls", "derived", "surfaces", "hsa_cmd", "wp_index")
    for x, y, name in zip(outs[0], outs[1], names):
        assert np.array_equal(x.view(np.uint32), y.view(np.uint32)), f"{level}: {name} differs in {(x != y).sum()} places"
    assert np.isfinite(outs[1][0]).all()
    if level == "waypoint" and not loop:
        done = outs[1][7] >= 2
        assert 0.2 < done.mean() < 1.0 or done.all()          # some aircraft finished (and at different steps)
        assert len(np.unique(outs[1][0][12][done])) > 3        # ... they stopped at different times
