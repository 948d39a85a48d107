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
    alive = np.ones(n, bool); worst = 0.0
    for t in range(steps):
        fleet.step_hsa(cmd[t // hold], dt, surfaces_out=surf)
        if (t + 1) % every:
            continue
        k = (t + 1) // every - 1
        gs = g["states"][k]
        alive &= ~((-gs[2] < 5.0) | (np.abs(gs[6]) > np.radians(120)) | (np.abs(gs[7]) > np.radians(80)))
        e = np.maximum(rel_err_cols(fleet.state.cpu().numpy(), gs), rel_err_cols(surf.cpu().numpy(), g["surfaces"][k]))
        worst = max(worst, float(e[alive].max()))
        assert (e[alive] <= TOL).all(), f"hsa step {t + 1}: {e[alive].max():.3e}"
    assert alive.sum() >= n // 2
    print(f"hsa_agent golden: worst rel err {worst:.3e}, {int(alive.sum())}/{n} in envelope")


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
    worst = 0.0
    for t in range(steps):
        fleet.step_waypoint(dt, surfaces_out=surf, hsa_cmd_out=hc)
        if (t + 1) % every:
            continue
        k = (t + 1) // every - 1
        assert np.array_equal(fleet.wp_index.cpu().numpy(), g["wp_index"][k]), f"{name}: waypoint index at step {t + 1}"
        e = max(rel_err(fleet.state.cpu().numpy(), g["states"][k]), rel_err(surf.cpu().numpy(), g["surfaces"][k]),
                rel_err(hc[0].cpu().numpy(), g["heading_cmd"][k]))
        worst = max(worst, e)
        assert e <= TOL, f"{name} step {t + 1}: {e:.3e}"
    print(f"{name} golden: worst rel err {worst:.3e} over {steps} steps; waypoint indices {fleet.wp_index.tolist()}")


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
    assert np.array_equal(out[32][1], out[64][1])
    e = rel_err_cols(out[32][0][:12], out[64][0][:12])
    assert e.max() <= 5e-3, e.max()          # 30 s of closed-loop flight
    print(f"fp32 vs fp64 after 3000 control steps of the square mission: worst rel err {e.max():.2e}")
