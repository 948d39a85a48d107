"""No kernel of the env path writes outside the caller's buffers.  Every buffer the C ABI is handed (fleet state, PID memory,
env bookkeeping, observations, flags, statistics) is re-homed into the middle of a larger allocation whose 512 bytes on either
side hold a sentinel pattern; after resets, whole-fleet steps, ranged steps with odd counts and host-buffer steps on ragged
fleets (pair kernel with a partial last block, odd fleets on the one-aircraft kernel, both builds) the bands must be intact.
(compute-sanitizer is not available on this pool; this is the bounds check the suite carries itself.)"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

PAD = 512
SENTINEL = 0xA5


def _rehome(t):
    """Copy tensor t into the middle of a sentinel-filled byte buffer; returns (view with t's shape/dtype, the byte buffer)."""
    nbytes = t.numel() * t.element_size()
    raw = torch.full((nbytes + 2 * PAD,), SENTINEL, dtype=torch.uint8, device=t.device)
    view = raw[PAD:PAD + nbytes].view(t.dtype).view(t.shape)
    view.copy_(t)
    return view, raw


def _guarded_env(acb200, n, controller, precision, **kw):
    env = acb200.BatchedRateControlEnv(n, controller=controller, precision=precision, seed=3, episode_length=0.2,
                                       track_components=True, **kw)
    from acb200 import _cabi
    raws = {}
    for name in ["state", "controls", "derived", "pid"] + list(_cabi.ENV_POINTER_FIELDS) + ["_scratch"]:
        t = getattr(env, name)
        if t is None:
            continue
        view, raw = _rehome(t)
        setattr(env, name, view)
        raws[name] = raw
        if name in ("state", "controls", "derived", "pid"):
            setattr(env.c.fleet, name, view.data_ptr())
        elif name != "_scratch":
            setattr(env.c, name, view.data_ptr())
    return env, raws


def _bands_intact(raws):
    torch.cuda.synchronize()
    for name, raw in raws.items():
        assert bool((raw[:PAD] == SENTINEL).all()), f"{name}: bytes in front of the buffer were written"
        assert bool((raw[-PAD:] == SENTINEL).all()), f"{name}: bytes behind the buffer were written"


@pytest.mark.parametrize("precision,controller,n", [(32, "attitude_pid", 2 * 1037), (32, "residual", 777), (32, "rate_pid", 2),
                                                    (32, "external", 130), (64, "attitude_pid", 1001), (64, "external", 129)])
def test_env_kernels_stay_inside_their_buffers(precision, controller, n):
    import acb200
    env, raws = _guarded_env(acb200, n, controller, precision)
    needs_action = controller in ("external", "residual")
    g = torch.Generator().manual_seed(1)
    env.reset()
    _bands_intact(raws)
    for t in range(25):                                   # 10-step episodes: auto-resets inside the step
        action = ((torch.rand(n, 4, generator=g) * 2 - 1) * 0.5).cuda() if needs_action else None
        env.step(action)
    _bands_intact(raws)
    mask = (torch.arange(n) % 3 == 0).to(torch.uint8)
    env.reset(mask)
    _bands_intact(raws)
    # ranged steps: the last range ends at n, counts odd and even
    env._bind_stream()
    action = ((torch.rand(n, 4, generator=g) * 2 - 1) * 0.5).cuda() if needs_action else None
    for first in range(0, n, 256):                        # odd counts; the last range ends at n
        count = min(n - first, 255)
        env.ctx.check(env.ctx.lib.acb200_env_step_range(env.ctx.handle, C.byref(env.c), C.byref(env.cfg),
                                                        None if action is None else C.c_void_p(action.data_ptr()),
                                                        first, count))
    _bands_intact(raws)
    assert torch.isfinite(env.state).all() and torch.isfinite(env.obs).all()


@pytest.mark.parametrize("precision,n", [(32, 2 * 1037), (32, 1001), (64, 515)])
def test_host_step_stays_inside_its_buffers(precision, n):
    import acb200
    env, raws = _guarded_env(acb200, n, "attitude_pid", precision)
    env.reset()
    host = {}
    for name, shape, dt in (("obs", (n, 18), torch.float32), ("reward", (n,), torch.float32),
                            ("terminated", (n,), torch.uint8), ("truncated", (n,), torch.uint8)):
        nbytes = int(torch.empty(shape, dtype=dt).numel()) * torch.empty((), dtype=dt).element_size()
        raw = torch.full((nbytes + 2 * PAD,), SENTINEL, dtype=torch.uint8).pin_memory()
        host[name] = (raw[PAD:PAD + nbytes].view(dt).view(shape), raw)
    att = torch.zeros((4, n), dtype=env.dtype).pin_memory()
    att[2] = float("inf"); att[3] = 0.6
    for chunks in (0, 3):
        env.ctx.check(env.ctx.lib.acb200_set_host_chunks(env.ctx.handle, chunks))
        for _ in range(12):
            env.step_host(None, host["obs"][0], host["reward"][0], host["terminated"][0], host["truncated"][0], att_cmd_host=att)
    _bands_intact(raws)
    for name, (view, raw) in host.items():
        assert bool((raw[:PAD] == SENTINEL).all()) and bool((raw[-PAD:] == SENTINEL).all()), f"host {name}: guard band written"
    assert torch.equal(host["obs"][0], env.obs.cpu()) and torch.isfinite(host["obs"][0]).all()


@pytest.mark.parametrize("level,n,dt", [("attitude", 2 * 517, 0.01), ("waypoint", 2 * 517, 0.01), ("waypoint", 2 * 517, 0.001),
                                        ("surface", 333, 0.01)])
def test_fleet_kernels_stay_inside_their_buffers(level, n, dt):
    import acb200
    fleet = acb200.AircraftFleet(n, precision=32)
    raws = {}
    for name in ("state", "controls", "derived", "pid", "pid_outer", "wp_index"):
        t = getattr(fleet, name, None)
        if t is None:
            continue
        view, raw = _rehome(t)
        setattr(fleet, name, view)
        raws[name] = raw
        if name in ("state", "controls", "derived", "pid"):
            setattr(fleet.c, name, view.data_ptr())
    init = torch.zeros(13, n)
    init[2] = -100.0; init[3] = 15.0
    fleet.reset(init.cuda())
    cmd, raws["cmd"] = _rehome(torch.zeros(4, n, device="cuda"))
    cmd[3] = 0.6
    surf, raws["surf"] = _rehome(torch.zeros(4, n, device="cuda"))
    if level == "waypoint":
        fleet.set_mission([[25.0, 0.0, -100.0, 15.0], [50.0, 20.0, -100.0, 15.0]], acceptance_radius=12.0, guidance="PP", loop=True)
    for _ in range(40):
        if level == "surface":
            fleet.step_surface(cmd, dt)
        elif level == "attitude":
            fleet.step_attitude(cmd, dt, surfaces_out=surf)
        else:
            fleet.step_waypoint(dt, surfaces_out=surf)
    _bands_intact(raws)
    assert torch.isfinite(fleet.state).all()


@pytest.mark.parametrize("kind,n", [("lstm", 300), ("lstm", 129), ("mlp", 77)])
def test_policy_forward_stays_inside_its_buffers(kind, n):
    """The tensor-core policy forward works on 128-aircraft tiles (TMA loads, tensor-memory epilogue): a ragged last tile
    must neither write past the action rows nor past the recurrent state of the aircraft that exist."""
    import acb200
    from acb200.policy import random_policy_state_dict
    pol = acb200.LearnedPolicyB200(random_policy_state_dict(kind, seed=0), kind=kind, n=n)
    raws = {}
    for name in ("h_state", "c_state", "action"):
        t = getattr(pol, name, None)
        if t is None:
            continue
        view, raw = _rehome(t)
        setattr(pol, name, view)
        raws[name] = raw
    g = torch.Generator().manual_seed(2)
    obs, raws["obs"] = _rehome(torch.randn(n, 18, generator=g).cuda())
    out, raws["out"] = _rehome(torch.zeros(n, 4, device="cuda"))
    for t in range(3):
        start = (torch.rand(n, generator=g) < (1.0 if t == 0 else 0.2)).to(torch.uint8).cuda()
        pol.forward(obs, start, out=out)
        pol.forward(obs, start)                       # into the policy's own action buffer
    _bands_intact(raws)
    assert torch.isfinite(out).all() and float(out.abs().max()) <= 1.0 + 1e-6


@pytest.mark.parametrize("precision,n", [(32, 1001), (64, 333)])
def test_sensor_update_stays_inside_its_buffers(precision, n):
    import acb200
    fleet = acb200.AircraftFleet(n, precision=precision)
    cfg = {"enabled": True, "imu_gyro_stddev": 0.02, "gps_position_stddev": 2.5, "attitude_stddev": 0.05, "seed": 11}
    sensor = acb200.NoisySensorB200(cfg, n, precision=precision)
    raws = {}
    for name in ("bias", "measured_state", "measured_derived"):
        view, raw = _rehome(getattr(sensor, name))
        setattr(sensor, name, view)
        raws[name] = raw
    for _ in range(20):
        sensor.update(fleet)
    sensor.reset((torch.arange(n) % 2 == 0).to(torch.uint8))
    sensor.update(fleet)
    _bands_intact(raws)
    assert torch.isfinite(sensor.measured_state).all()


@pytest.mark.parametrize("T,n", [(7, 1001), (33, 130)])
def test_gae_stays_inside_its_buffers(T, n):
    import acb200
    buf = acb200.RolloutBufferB200(T, n)
    raws = {}
    for name in ("rewards", "values", "episode_starts", "advantages", "returns"):
        view, raw = _rehome(getattr(buf, name))
        setattr(buf, name, view)
        raws[name] = raw
    g = torch.Generator().manual_seed(4)
    buf.rewards.copy_(torch.randn(T, n, generator=g)); buf.values.copy_(torch.randn(T, n, generator=g))
    buf.episode_starts.copy_((torch.rand(T, n, generator=g) < 0.1).to(torch.uint8))
    buf.pos, buf.full = T, True
    buf.compute_returns_and_advantage(torch.randn(n, generator=g).cuda(), (torch.rand(n, generator=g) < 0.1).cuda())
    _bands_intact(raws)
    assert torch.isfinite(buf.advantages).all() and torch.isfinite(buf.returns).all()


@pytest.mark.parametrize("level", ["hsa", "waypoint"])
@pytest.mark.parametrize("n,dtp", [(4, 0.01), (132, 0.01), (1000, 0.01), (4100, 0.01), (1002, 0.01), (1000, 0.001)])
def test_guided_fleet_steps_stay_inside_their_buffers(level, n, dtp):
    """Levels 2 and 1 on a fleet: K = 1 with the fleet a multiple of 4 takes the TMA-staged kernel (tensor stores of 128-aircraft
    boxes: the engine must clip the ragged last block at n), 1002 the direct one-aircraft kernel, K = 10 the pair kernel."""
    import acb200
    fleet = acb200.AircraftFleet(n, precision=32, config={"dt_physics": dtp})
    raws = {}
    for name in ("state", "controls", "derived", "pid", "pid_outer", "wp_index"):
        view, raw = _rehome(getattr(fleet, name))
        setattr(fleet, name, view)
        raws[name] = raw
        if name in ("state", "controls", "derived", "pid"):
            setattr(fleet.c, name, view.data_ptr())
    surf, raws["surfaces_out"] = _rehome(torch.zeros(4, n, device="cuda"))
    hc, raws["hsa_cmd_out"] = _rehome(torch.zeros(3, n, device="cuda"))
    g = torch.Generator().manual_seed(2)
    init = torch.zeros(13, n)
    init[0].uniform_(-20, 20, generator=g); init[1].uniform_(-20, 20, generator=g); init[2] = -100.0; init[3] = 15.0
    fleet.reset(init.cuda())
    _bands_intact(raws)
    if level == "waypoint":
        fleet.set_mission([[25.0, 0.0, -100.0, 15.0], [50.0, 20.0, -100.0, 15.0]], acceptance_radius=12.0, guidance="PP", loop=False)
        assert fleet._mission.wp_index == fleet.wp_index.data_ptr()
    cmd = torch.zeros(3, n, device="cuda"); cmd[1] = 15.0; cmd[2] = 100.0
    for _ in range(60):
        if level == "hsa":
            fleet.step_hsa(cmd, 0.01, surfaces_out=surf)
        else:
            fleet.step_waypoint(0.01, surfaces_out=surf, hsa_cmd_out=hc)
    _bands_intact(raws)
    assert bool(torch.isfinite(fleet.state).all()) and float(fleet.state[12].max()) > 0.5      # the fleet really flew
