"""Levels 2 (HSAAgent / TECS) and 1 (MissionPlanner + WaypointAgent) of the cascade on the GPU: fp64 build against the
golden trajectories generated from the reference and against the CPU oracle; fp32 build against the fp64 build."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, rel_err
from helpers import rel_err_cols

pytestmark = pytest.mark.gpu
TOL = 1e-9


@pytest.fixture(scope="module")
def mods():
    import torch
    import acb200
    from oracle import oracle as orc
    return torch, acb200, orc


def _t(torch, a, dtype=None):
    return torch.as_tensor(np.ascontiguousarray(a), device="cuda", dtype=dtype)


def test_hsa_golden(mods):
    torch, acb200, orc = mods
    g = np.load(os.path.join(GOLDEN, "hsa_agent.npz"))
    n = g["s0"].shape[1]
    hold, steps, dt, every = int(g["hold"]), int(g["steps"]), float(g["dt"]), int(g["every"])
    fleet = acb200.AircraftFleet(n, precision=64)
    fleet.reset(_t(torch, g["s0"]))
    cmd = _t(torch, g["cmd"]); surf = torch.zeros((4, n), device="cuda", dtype=torch.float64)
    alive = np.ones(n, bool); worst = 0.0; events = 0
    for t in range(steps):
        fleet.step_hsa(cmd[t // hold], dt, surfaces_out=surf)
        if (t + 1) % every:
            continue
        k = (t + 1) // every - 1
        gs = g["states"][k]
        alive &= ~((-gs[2] < 5.0) | (np.abs(gs[6]) > np.radians(120)) | (np.abs(gs[7]) > np.radians(80)))
        e = np.maximum(rel_err_cols(fleet.state.cpu().numpy(), gs), rel_err_cols(surf.cpu().numpy(), g["surfaces"][k]))
        # DISCONTINUITY EVENTS: the heading error is wrapped to [-pi, pi) (hsa_agent.py:147); when the aircraft passes
        # exactly opposite to the commanded heading a 1e-14 difference between two libms decides whether it keeps turning
        # left or right.  Such an aircraft jumps from ~1e-13 to ~1e-6 within one snapshot (the CPU oracle reproduces the
        # golden file through these events; it shares the reference's libm).  It leaves the comparison and is counted.
        jumped = alive & (e > TOL)
        events += int(jumped.sum()); alive &= ~jumped
        worst = max(worst, float(e[alive].max()))
    assert events <= n // 4, events
    assert alive.sum() >= n // 2
    print(f"hsa_agent golden: worst rel err {worst:.3e} over {steps} steps, {int(alive.sum())}/{n} compared to the end, "
          f"{events} heading-wrap discontinuity event(s)")


@pytest.mark.parametrize("name", ["waypoint_pp", "waypoint_los"])
def test_waypoint_mission_golden(mods, name):
    torch, acb200, orc = mods
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    n = g["s0"].shape[1]
    steps, dt, every = int(g["steps"]), float(g["dt"]), int(g["every"])
    fleet = acb200.AircraftFleet(n, precision=64, config={"dt_physics": dt})      # the demo steps Simplified6DOF directly: K = 1
    fleet.reset(_t(torch, g["s0"]))
    fleet.set_mission(g["waypoints"], acceptance_radius=float(g["acceptance_radius"]), guidance="LOS" if int(g["guidance"]) == 0 else "PP")
    surf = torch.zeros((4, n), device="cuda", dtype=torch.float64); hc = torch.zeros((3, n), device="cuda", dtype=torch.float64)
    worst = 0.0; ok = np.ones(n, bool); events = 0
    for t in range(steps):
        fleet.step_waypoint(dt, surfaces_out=surf, hsa_cmd_out=hc)
        if (t + 1) % every:
            continue
        k = (t + 1) // every - 1
        e = np.maximum.reduce([rel_err_cols(fleet.state.cpu().numpy(), g["states"][k]), rel_err_cols(surf.cpu().numpy(), g["surfaces"][k]),
                               np.abs(hc[0].cpu().numpy() - g["heading_cmd"][k])])
        # same discontinuity events as in the HSA test, plus the LOS turn-anticipation switch (waypoint_agent.py:139-145)
        jumped = ok & (e > TOL)
        events += int(jumped.sum()); ok &= ~jumped
        worst = max(worst, float(e[ok].max()))
        assert np.array_equal(fleet.wp_index.cpu().numpy()[ok], g["wp_index"][k][ok]), f"{name}: waypoint index at step {t + 1}"
    assert events <= n // 4, events
    print(f"{name} golden: worst rel err {worst:.3e} over {steps} steps, {int(ok.sum())}/{n} compared to the end, {events} "
          f"discontinuity event(s); waypoint indices {fleet.wp_index.tolist()}")


def test_mission_vs_oracle_many_aircraft_loop_and_completion(mods):
    """C5 shape at reduced N: jittered starts, looping mission on the GPU vs the oracle; then a finite mission completes
    and completed aircraft stop moving (the demo's `break`)."""
    torch, acb200, orc = mods
    n, steps, dt = 1024, 1500, 0.01
    rng = np.random.default_rng(8)
    size, alt, spd = 150.0, 100.0, 15.0
    wps = np.array([[size, 0, -alt, spd], [size, size, -alt, np.nan], [0, size, -alt, spd], [0, 0, -alt, spd]])
    s0 = np.zeros((13, n)); s0[2] = -alt + rng.uniform(-5, 5, n); s0[3] = spd
    s0[0] = rng.uniform(-20, 20, n); s0[1] = rng.uniform(-20, 20, n); s0[8] = rng.uniform(-np.radians(10), np.radians(10), n)
    P, Cp = orc.default_aircraft_params(), orc.default_controller_params()
    for guidance, gid in (("PP", 1), ("LOS", 0)):
        fleet = acb200.AircraftFleet(n, precision=64, config={"dt_physics": dt})
        fleet.reset(_t(torch, s0)); fleet.set_mission(wps, acceptance_radius=40.0, guidance=guidance, loop=True)
        ref = s0.copy(); idx = np.zeros(n, np.int32); pid = np.zeros((18, n), np.float32); outer = np.zeros((9, n), np.float32)
        ok = np.ones(n, bool); worst = 0.0
        for t in range(steps):
            fleet.step_waypoint(dt)
            orc.step_waypoint(P, Cp, ref, wps, idx, pid, outer, dt, guidance=gid, acceptance_radius=40.0, loop=True, dt_physics=dt)
            if (t + 1) % 50 == 0:
                same = (np.all(fleet.pid.cpu().numpy().view(np.uint32) == pid.view(np.uint32), axis=0)
                        & np.all(fleet.pid_outer.cpu().numpy().view(np.uint32) == outer.view(np.uint32), axis=0)
                        & (fleet.wp_index.cpu().numpy() == idx))
                ok &= same                     # a float rounding flip at a PID input ends bit-comparability for that aircraft
                e = rel_err_cols(fleet.state.cpu().numpy(), ref)
                worst = max(worst, float(e[ok].max()))
                assert (e[ok] <= TOL).all(), f"{guidance} step {t + 1}: {e[ok].max():.3e}"
        assert ok.mean() > 0.95, ok.mean()
        assert idx.max() >= 1
        print(f"mission {guidance} vs oracle: worst rel err {worst:.3e}, {ok.mean():.1%} bit-comparable to the end, max waypoint index {idx.max()}")
    # finite mission: everybody finishes the short leg, then stays put
    fleet = acb200.AircraftFleet(64, precision=64, config={"dt_physics": dt})
    fleet.set_mission(np.array([[0.0, 0.0, -100.0, 15.0], [100.0, 0.0, -100.0, 15.0]]), acceptance_radius=40.0)
    for _ in range(800):
        fleet.step_waypoint(dt)
    assert bool(fleet.mission_complete().all())
    frozen = fleet.state.clone()
    fleet.step_waypoint(dt)
    assert torch.equal(frozen, fleet.state)


def test_fp32_build_flies_the_same_mission(mods):
    torch, acb200, orc = mods
    g = np.load(os.path.join(GOLDEN, "waypoint_pp.npz"))
    n = g["s0"].shape[1]; dt = float(g["dt"])
    out = {}
    for precision in (64, 32):
        fleet = acb200.AircraftFleet(n, precision=precision, config={"dt_physics": dt})
        fleet.reset(_t(torch, g["s0"])); fleet.set_mission(g["waypoints"], acceptance_radius=40.0, guidance="PP")
        for t in range(3000):
            fleet.step_waypoint(dt)
        out[precision] = (fleet.state.double().cpu().numpy(), fleet.wp_index.cpu().numpy())
    assert np.array_equal(out[32][1], out[64][1])          # same waypoint reached by every aircraft
    a, b = out[32][0], out[64][0]
    dpos = np.hypot(a[0] - b[0], a[1] - b[1]); dalt = np.abs(a[2] - b[2]); dV = np.abs(np.linalg.norm(a[3:6], axis=0) - np.linalg.norm(b[3:6], axis=0))
    print(f"fp32 vs fp64 after 3000 control steps (30 s) of the square mission: ground track {dpos.max():.3f} m, altitude {dalt.max():.4f} m, "
          f"airspeed {dV.max():.4f} m/s apart at worst")
    assert dpos.max() <= 5.0 and dalt.max() <= 0.5 and dV.max() <= 0.2


@pytest.mark.parametrize("K,loop", [(1, True), (10, False)])
def test_mission_statistics_match_the_oracle(K, loop):
    """acb200_mission.stats (BASELINE config 5's fleet statistics) from the one-aircraft kernel (K = 1) and the pair kernel
    (K = 10, fp32) against the oracle's: event counts identical (fp64), sums to round-off / the fp32 tolerance."""
    import torch
    import acb200
    from oracle import oracle as orc
    from tools.bench_c5 import SQUARE, initial_states
    n, steps, dt = 1024, 400, 0.01
    dtp = dt / K
    s0, idx0 = initial_states(n, 0)
    wps = np.asarray(SQUARE, np.float64)
    P, CP = orc.default_aircraft_params(), orc.default_controller_params()
    state = s0.copy(); idx = idx0.copy()        # copies: the oracle advances them in place
    pid = np.zeros((18, n), np.float32); outer = np.zeros((9, n), np.float32); ref = np.zeros(16)
    for _ in range(steps):
        orc.step_waypoint(P, CP, state, wps, idx, pid, outer, dt, guidance=orc.GUIDANCE_PP, acceptance_radius=40.0, loop=loop,
                          dt_physics=dtp, mission_stats=ref)
    assert ref[1] > n // 20 and ref[0] > 0.9 * n * steps                 # the advance branch ran
    if not loop:
        assert ref[3] > 0                                                # some aircraft completed the open mission
    for precision, rtol in ((64, 1e-9), (32, 2e-3)):
        fleet = acb200.AircraftFleet(n, precision=precision, config={"dt_physics": dtp})
        fleet.reset(torch.as_tensor(s0))
        fleet.set_mission(SQUARE, acceptance_radius=40.0, guidance="PP", loop=loop)
        fleet.wp_index.copy_(torch.as_tensor(idx0))
        for _ in range(steps):
            fleet.step_waypoint(dt)
        got = fleet.mission_stats_buf.cpu().numpy()
        if precision == 64:
            assert np.array_equal(got[[0, 1, 2, 3, 9]], ref[[0, 1, 2, 3, 9]]), (got, ref)
            assert np.array_equal(fleet.wp_index.cpu().numpy(), idx)
        else:
            assert np.allclose(got[[0, 1, 2, 3]], ref[[0, 1, 2, 3]], rtol=5e-3), (got, ref)
        assert np.allclose(got[[4, 5, 6, 7, 8]], ref[[4, 5, 6, 7, 8]], rtol=rtol), (precision, got, ref)
        st = fleet.mission_stats(reset=True)
        assert st["waypoints_reached"] == got[1] and abs(st["airspeed_mean"] - got[8] / got[0]) < 1e-12
        assert float(fleet.mission_stats_buf.abs().sum()) == 0.0
    # stats=False: no accumulator, same trajectories
    a = acb200.AircraftFleet(n, precision=32, config={"dt_physics": dtp}); b = acb200.AircraftFleet(n, precision=32, config={"dt_physics": dtp})
    for f, st in ((a, True), (b, False)):
        f.reset(torch.as_tensor(s0)); f.set_mission(SQUARE, guidance="PP", loop=loop, stats=st); f.wp_index.copy_(torch.as_tensor(idx0))
        for _ in range(50):
            f.step_waypoint(dt)
    assert torch.equal(a.state, b.state) and b.mission_stats_buf is None


# fp32 build, ONE control step from the oracle's state (re-synchronised every step, controller memory included): what the
# throughput build changes in the guidance levels -- angle wraps without fmod / sincos+atan2, MUFU quotients, the pure-pursuit
# bearing taken as the waypoint bearing, the SFU cosine of the bank angle, the fused PID division -- is compared here with
# the fp64 oracle on the command it issues, the surfaces, the controller memory and the state after the step.
# Bars (x max(1, |value|)); measured worst on B200 in the comment.  Aircraft sitting on a discontinuity of the guidance law
# (waypoint switch, 20 deg / turn-threshold tests: the oracle's own output moves by > 1e-4 when its state moves by one float
# ulp) are excluded and counted.
GUIDE_CMD_TOL = 4e-6        # heading (on the circle), speed, altitude command -- what the guidance shortcuts produce; measured 8.5e-7
GUIDE_SURF_TOL = 1e-3       # elevator, aileron, rudder, throttle; measured 2.4e-4.  From here on the error is fp32 conditioning of the
                            # cascade, not guidance: the TECS energies g h + V^2 / 2 are O(1e3), one float ulp of them is 1.2e-4, and
                            # every PID level multiplies an input difference by ~10 through its derivative term (kd x 0.1 / dt)
GUIDE_PID_TOL = 1.5e-3      # integral and previous-error words of the nine PIDs; measured 3.5e-4
GUIDE_PID_DER_TOL = 8e-3    # their filtered-derivative words (the pitch-rate loop, three levels down, is the worst); measured 1.8e-3
GUIDE_STATE_TOL = 1.5e-4    # position, velocity, attitude after the step (open loop, same surfaces: 5e-6 in test_gpu_parity_f32.py);
                            # measured 3.9e-5
GUIDE_RATE_TOL = 8e-4       # body rates after the step: the surfaces' difference x control power (up to 360 rad/s^2 per unit of
                            # aileron at 30 m/s) x dt; measured 1.9e-4


@pytest.mark.parametrize("guidance,gid", [("PP", 1), ("LOS", 0)])
@pytest.mark.parametrize("dtp", [0.01, 0.001])                     # K = 1: one-aircraft kernel; K = 10: pair kernel
def test_fp32_guidance_per_step_from_resynchronised_states(mods, guidance, gid, dtp):
    torch, acb200, orc = mods
    n, steps, dt, R = 4096, 120, 0.01, 40.0
    rng = np.random.default_rng(21)
    wps = np.array([[150.0, 0, -100.0, 15.0], [150.0, 150.0, -100.0, np.nan], [0, 150.0, -100.0, 18.0], [0, 0, -100.0, 15.0]])
    ref = np.zeros((13, n))
    ref[0] = rng.uniform(-100, 250, n); ref[1] = rng.uniform(-100, 250, n); ref[2] = -100 + rng.uniform(-10, 10, n)
    ref[3] = rng.uniform(12, 25, n); ref[4] = rng.uniform(-1, 1, n); ref[5] = rng.uniform(-1, 1, n)
    ref[6] = rng.uniform(-0.3, 0.3, n); ref[7] = rng.uniform(-0.15, 0.15, n); ref[8] = rng.uniform(-np.pi, np.pi, n)
    ref[9:12] = rng.uniform(-0.2, 0.2, (3, n))
    idx = rng.integers(0, 4, n).astype(np.int32)
    pid = np.zeros((18, n), np.float32); outer = np.zeros((9, n), np.float32)
    P, Cp = orc.default_aircraft_params(), orc.default_controller_params()
    fleet = acb200.AircraftFleet(n, precision=32, config={"dt_physics": dtp})
    fleet.set_mission(wps, acceptance_radius=R, guidance=guidance, loop=True, stats=False)
    surf = torch.zeros(4, n, device="cuda"); hc = torch.zeros(3, n, device="cuda")
    rs, rh, rs2, rh2 = (np.zeros((k, n)) for k in (4, 3, 4, 3))
    worst = np.zeros(6); excluded = 0; switched = 0

    def rel(a, b):
        return np.abs(a - b) / np.maximum(np.abs(b), 1.0)

    for t in range(steps):
        start32 = ref.astype(np.float32)
        fleet.reset(_t(torch, start32))                       # state + derived; zeroes the controller memory ...
        fleet.pid.copy_(_t(torch, pid)); fleet.pid_outer.copy_(_t(torch, outer)); fleet.wp_index.copy_(_t(torch, idx))   # ... so it goes in again
        fleet.step_waypoint(dt, surfaces_out=surf, hsa_cmd_out=hc)
        # the oracle's sensitivity: the same step from a state one float ulp away
        pert = np.nextafter(start32, np.float32(np.inf)).astype(np.float64)
        i2, p2, o2 = idx.copy(), pid.copy(), outer.copy()
        orc.step_waypoint(P, Cp, pert, wps, i2, p2, o2, dt, guidance=gid, acceptance_radius=R, loop=True, dt_physics=dtp,
                          surfaces_out=rs2, hsa_cmd_out=rh2)
        ref = start32.astype(np.float64)
        idx_before = idx.copy()
        orc.step_waypoint(P, Cp, ref, wps, idx, pid, outer, dt, guidance=gid, acceptance_radius=R, loop=True, dt_physics=dtp,
                          surfaces_out=rs, hsa_cmd_out=rh)
        switched += int((idx != idx_before).sum())
        dh2 = np.abs(np.angle(np.exp(1j * (rh2[0] - rh[0]))))
        smooth = (i2 == idx) & (dh2 < 1e-4) & (np.abs(rh2[1] - rh[1]) < 1e-4) & (np.abs(rs2 - rs).max(axis=0) < 1e-3)
        excluded += int((~smooth).sum())
        got_h, got_s = hc.double().cpu().numpy(), surf.double().cpu().numpy()
        assert np.array_equal(fleet.wp_index.cpu().numpy()[smooth], idx[smooth]), f"step {t + 1}: waypoint index"
        e_cmd = np.maximum(np.abs(np.angle(np.exp(1j * (got_h[0] - rh[0])))), rel(got_h[1:], rh[1:]).max(axis=0))
        e_surf = rel(got_s, rs).max(axis=0)
        e_pid_rows = np.concatenate([rel(fleet.pid.double().cpu().numpy(), pid.astype(np.float64)),
                                     rel(fleet.pid_outer.double().cpu().numpy(), outer.astype(np.float64))])
        der = np.arange(27) % 3 == 2
        e_pid, e_der = e_pid_rows[~der].max(axis=0), e_pid_rows[der].max(axis=0)
        got_state = fleet.state.double().cpu().numpy()
        e_state, e_rate = rel_err_cols(got_state[:9], ref[:9]), rel_err_cols(got_state[9:12], ref[9:12])
        for k, (e, tol, name) in enumerate(((e_cmd, GUIDE_CMD_TOL, "HSA command"), (e_surf, GUIDE_SURF_TOL, "surfaces"),
                                           (e_pid, GUIDE_PID_TOL, "controller memory"), (e_state, GUIDE_STATE_TOL, "state"),
                                           (e_der, GUIDE_PID_DER_TOL, "filtered derivatives"), (e_rate, GUIDE_RATE_TOL, "body rates"))):
            worst[k] = max(worst[k], float(e[smooth].max()))
            assert (e[smooth] <= tol).all(), f"{guidance} dt_physics={dtp} step {t + 1}: {name} {e[smooth].max():.3e} > {tol}"
    print(f"fp32 guidance per step ({guidance}, dt_physics={dtp}): command {worst[0]:.2e}, surfaces {worst[1]:.2e}, controller memory "
          f"{worst[2]:.2e} / derivatives {worst[4]:.2e}, state {worst[3]:.2e} / rates {worst[5]:.2e}; {excluded} of {n * steps} aircraft-steps on a discontinuity, {switched} waypoint switches")
    assert excluded <= 0.01 * n * steps and switched > 0


@pytest.mark.parametrize("level", ["hsa", "waypoint"])
@pytest.mark.parametrize("n", [4, 128, 132, 1000, 4100])
def test_tma_staged_guided_step_is_bit_identical_to_the_direct_kernel(mods, level, n):
    """fp32, K = 1: the guided step whose state / derived / controller-memory words travel through shared memory as 2-D TMA boxes
    (step_guided_staged_kernel) against the kernel that addresses the fleet directly (ACB200_GUIDED_STAGED=0), every buffer
    bit for bit -- whole blocks, a ragged last block that the TMA engine clips (132, 1000, 4100), a fleet smaller than a
    block, and a finite mission in which aircraft finish at different steps (their columns go back as loaded)."""
    torch, acb200, orc = mods
    dt, steps = 0.01, 400
    g = torch.Generator().manual_seed(5)
    init = torch.zeros(13, n)
    init[0].uniform_(-20, 20, generator=g); init[1].uniform_(-20, 20, generator=g); init[2] = -100.0
    init[3] = 15.0; init[8].uniform_(-0.2, 0.2, generator=g)
    cmd = torch.zeros(3, n)
    cmd[0].uniform_(-1.0, 1.0, generator=g); cmd[1].uniform_(13, 18, generator=g); cmd[2].uniform_(80, 120, generator=g)
    wps = [[25.0, 0.0, -100.0, 15.0], [50.0, 20.0, -100.0, 15.0]]
    outs = []
    for staged in ("1", "0"):
        os.environ["ACB200_GUIDED_STAGED"] = staged
        try:
            fleet = acb200.AircraftFleet(n, precision=32, config={"dt_physics": dt})
            fleet.reset(init.cuda())
            surf = torch.zeros(4, n, device="cuda"); hc = torch.zeros(3, n, device="cuda")
            if level == "waypoint":
                fleet.set_mission(wps, acceptance_radius=12.0, guidance="PP", loop=False)
            c = cmd.cuda()
            for _ in range(steps):
                if level == "hsa":
                    fleet.step_hsa(c, dt, surfaces_out=surf)
                else:
                    fleet.step_waypoint(dt, surfaces_out=surf, hsa_cmd_out=hc)
            torch.cuda.synchronize()
            bufs = [fleet.state, fleet.pid, fleet.pid_outer, fleet.controls, fleet.derived, surf, hc]
            if level == "waypoint":
                bufs += [fleet.wp_index, fleet.mission_stats_buf]
            outs.append([t.cpu().numpy().copy() for t in bufs])
        finally:
            os.environ.pop("ACB200_GUIDED_STAGED", None)
    names = ("state", "pid", "pid_outer", "controls", "derived", "surfaces", "hsa_cmd", "wp_index", "mission_stats")
    for x, y, name in zip(outs[0], outs[1], names):
        if name == "mission_stats":               # double atomics across blocks: the order of the additions is not fixed
            assert np.allclose(x, y, rtol=1e-12, atol=0), name
        else:
            assert np.array_equal(x.view(np.uint8), y.view(np.uint8)), f"{level} n={n}: {name} differs in {(x != y).sum()} places"
    assert np.isfinite(outs[0][0]).all() and np.ptp(outs[0][0][0]) > 1.0
    if level == "waypoint" and n >= 128:
        assert (outs[0][7] >= 2).any() and len(np.unique(outs[0][0][12][outs[0][7] >= 2])) > 1      # finished, at different times


@pytest.mark.parametrize("level", ["hsa", "waypoint"])
@pytest.mark.parametrize("dtp", [0.01, 0.001])
def test_guided_step_without_the_derived_buffer_derives_the_same_values(mods, level, dtp):
    """include/acb200.h: the HSA / waypoint steps read airspeed, altitude and heading from acb200_fleet.derived when the fleet has
    that buffer and derive them from the state when it is NULL.  Both must give the same bits: the buffer holds what
    derived_outputs(state) returns (K = 1: TMA-staged kernel against the direct one; K = 10: the pair kernel in both cases)."""
    torch, acb200, orc = mods
    n, dt = 1024, 0.01
    g = torch.Generator().manual_seed(11)
    init = torch.zeros(13, n)
    init[0].uniform_(-20, 20, generator=g); init[1].uniform_(-20, 20, generator=g); init[2] = -100.0
    init[3] = 15.0; init[8].uniform_(-0.2, 0.2, generator=g)
    cmd = torch.zeros(3, n); cmd[0].uniform_(-1.0, 1.0, generator=g); cmd[1] = 15.0; cmd[2] = 100.0
    outs = []
    for with_derived in (True, False):
        fleet = acb200.AircraftFleet(n, precision=32, config={"dt_physics": dtp})
        fleet.reset(init.cuda())
        if not with_derived:
            fleet.c.derived = None
        if level == "waypoint":
            fleet.set_mission([[25.0, 0.0, -100.0, 15.0], [50.0, 20.0, -100.0, 15.0]], acceptance_radius=12.0, guidance="LOS", loop=True)
        c = cmd.cuda()
        for _ in range(200):
            fleet.step_hsa(c, dt) if level == "hsa" else fleet.step_waypoint(dt)
        torch.cuda.synchronize()
        outs.append([t.cpu().numpy().copy() for t in (fleet.state, fleet.pid, fleet.pid_outer, fleet.controls, fleet.wp_index)])
    for x, y, name in zip(outs[0], outs[1], ("state", "pid", "pid_outer", "controls", "wp_index")):
        assert np.array_equal(x.view(np.uint8), y.view(np.uint8)), f"{level} dt_physics={dtp}: {name} differs in {(x != y).sum()} places"
    assert np.ptp(outs[0][0][0]) > 1.0
