"""HSAAgent / WaypointAgent / MissionPlanner drop-in objects (acb200.guidance) and the YAML loader (acb200.config_loader).

Pinned on the golden closed-loop trajectories the unmodified reference agents flew (tests/golden/hsa_agent.npz,
waypoint_pp.npz, waypoint_los.npz): the objects drive the CPU restatement of the physics (oracle.step_surface, itself pinned on
the reference) with the reference's loop structure, and must land on the reference's states to 1e-9 -- with the waypoint index
sequence identical.  No GPU: the objects run on the library's host-side PID core."""
import os

import numpy as np
import pytest

import acb200
from acb200 import config_loader as cl
from acb200.adapters import AircraftState, ControlCommand, ControlMode, ControllerConfig, ControlSurfaces, Waypoint
from acb200.guidance import HSAAgent, MissionPlanner, MissionState, WaypointAgent
from conftest import GOLDEN, rel_err
from oracle import oracle as orc

TOL = 1e-9


def _state(s, time=None):
    d = orc.derived(np.ascontiguousarray(s[:12]))
    return AircraftState(time=float(s[12]) if time is None else time, position=s[0:3].copy(), velocity=s[3:6].copy(),
                         attitude=s[6:9].copy(), angular_rate=s[9:12].copy(), airspeed=float(d[0]), altitude=float(d[1]),
                         ground_speed=float(d[2]), heading=float(d[3]))


def test_hsa_agent_object_flies_the_reference_trajectory():
    g = np.load(os.path.join(GOLDEN, "hsa_agent.npz"))
    P = orc.default_aircraft_params()
    hold, steps, dt, every = int(g["hold"]), 1500, float(g["dt"]), int(g["every"])
    for i in (0, 3):
        agent = HSAAgent(ControllerConfig())
        assert agent.attitude_agent is not None and agent.get_control_level() == ControlMode.HSA
        s = np.ascontiguousarray(g["s0"][:, i:i + 1])
        for t in range(steps):
            c = g["cmd"][t // hold][:, i]
            surf = agent.compute_action(ControlCommand(mode=ControlMode.HSA, heading=c[0], speed=c[1], altitude=c[2]), _state(s[:, 0]), dt)
            orc.step_surface(P, s, np.ascontiguousarray(surf.to_array().reshape(4, 1)), dt)
            if (t + 1) % every == 0:
                k = (t + 1) // every - 1
                assert rel_err(s[:, 0], g["states"][k][:, i]) <= TOL, (i, t + 1)
                assert rel_err(surf.to_array(), g["surfaces"][k][:, i]) <= TOL, (i, t + 1)


@pytest.mark.parametrize("name", ["waypoint_pp", "waypoint_los"])
def test_waypoint_agent_and_mission_planner_objects_fly_the_reference_mission(name):
    """The loop of examples/03_waypoint_square_demo.py:147-209 with this package's objects."""
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    P = orc.default_aircraft_params()
    steps, dt, every = 4500, float(g["dt"]), int(g["every"])
    wps = [Waypoint(north=w[0], east=w[1], down=w[2], speed=None if np.isnan(w[3]) else float(w[3])) for w in g["waypoints"]]
    i = 2
    agent = WaypointAgent(ControllerConfig(), guidance_type="PP" if int(g["guidance"]) == 1 else "LOS")
    assert agent.hsa_agent is not None and agent.hsa_agent.attitude_agent is not None
    mission = MissionPlanner(wps, acceptance_radius=float(g["acceptance_radius"]))
    assert mission.state == MissionState.IDLE and mission.get_waypoint_command() is None
    mission.start()
    s = np.ascontiguousarray(g["s0"][:, i:i + 1])
    surf = ControlSurfaces()
    for t in range(steps):
        st = _state(s[:, 0])
        mission.update(st)
        if not mission.is_complete():
            surf = agent.compute_action(mission.get_waypoint_command(), st, dt)
            orc.step_surface(P, s, np.ascontiguousarray(surf.to_array().reshape(4, 1)), dt, dt_physics=dt)   # sim.step(dt): one RK4 step
        if (t + 1) % every == 0:
            k = (t + 1) // every - 1
            assert mission.current_waypoint_index == int(g["wp_index"][k][i]), (t + 1)
            assert rel_err(s[:, 0], g["states"][k][:, i]) <= TOL, (t + 1)
            assert abs(agent.last_heading_cmd - g["heading_cmd"][k][i]) <= 1e-9
    assert mission.current_waypoint_index >= 1 and mission.waypoints_reached == mission.current_waypoint_index
    assert mission.get_summary()["total_distance_m"] == pytest.approx(1200.0)
    assert np.array_equal(np.isnan(mission.waypoint_table()), np.isnan(g["waypoints"]))


def test_agent_argument_errors_match_the_reference():
    a = HSAAgent()
    with pytest.raises(AssertionError, match="HSA agent expects HSA mode"):
        a.compute_action(ControlCommand(mode=ControlMode.RATE), AircraftState())
    with pytest.raises(ValueError, match="'altitude' is None"):
        a.compute_action(ControlCommand(mode=ControlMode.HSA, heading=0.0, speed=15.0), AircraftState())
    with pytest.raises(AssertionError, match="Waypoint agent expects WAYPOINT mode"):
        WaypointAgent().compute_action(ControlCommand(mode=ControlMode.HSA), AircraftState())
    with pytest.raises(ValueError, match="at least one waypoint"):
        MissionPlanner([])
    m = MissionPlanner([Waypoint.from_altitude(10.0, 0.0, 100.0)], acceptance_radius=15.0)
    m.start()
    assert m.update(AircraftState(position=np.array([0.0, 0.0, -100.0]))) and m.is_complete() and m.get_progress_percentage() == 100.0
    m.abort(); assert m.is_aborted()


YAML = """
rate_control:
  roll: {kp: 2.0, ki: 0.5}
  limits: {max_roll_rate: 150.0}
hsa_control:
  heading: {kp: 1.5, ki: 0.1, kd: 0.3}
  tecs: {energy: {kp: 0.2, ki: 0.04, kd: 0.01}}
  limits: {max_bank_angle: 35.0, baseline_throttle: 0.2, max_pitch_command: 12.0}
  turn_compensation: {load_factor_gain: 0.08}
waypoint_navigation:
  acceptance_radius: 25.0
  pure_pursuit: {lookahead_time: 2.0}
  los: {max_bank_for_turn_calc: 15.0}
  speed_control: {min_speed: 12.0}
"""


def test_yaml_loader_and_controller_params(tmp_path):
    (tmp_path / "controllers").mkdir()
    (tmp_path / "controllers" / "custom.yaml").write_text(YAML)
    fc = cl.load_controller_config("custom.yaml", config_dir=tmp_path)
    assert fc.roll_rate_gains == cl.PIDGains(2.0, 0.5, 0.0) and fc.max_roll_rate == 150.0 and fc.max_pitch_rate == 100.0
    assert fc.hsa.max_bank_angle == 35.0 and fc.hsa.tecs.baseline_throttle == 0.2 and fc.hsa.tecs.load_factor_gain == 0.08
    assert fc.guidance.acceptance_radius == 25.0 and fc.guidance.lookahead_time == 2.0 and fc.guidance.lookahead_min == 12.0
    assert fc.guidance.los_max_bank == 15.0 and fc.guidance.min_speed == 12.0 and fc.guidance.pn_gain == 3.0
    assert cl.load_controller_config("missing.yaml", config_dir=tmp_path) == cl.FlightControlConfig()
    p = cl.controller_params(ControllerConfig(), fc)
    f32 = lambda v: float(np.float32(v))
    assert p.heading.kp == f32(1.5) and p.heading.output_max == f32(np.radians(35.0)) and p.heading.integral_max == 25.0
    assert p.energy.kp == f32(0.2) and p.energy.integral_max == 10.0 and p.balance.output_max == f32(np.radians(12.0))
    assert p.max_bank == np.radians(35.0) and p.max_bank_angle_deg == 35.0 and p.baseline_throttle == 0.2
    assert p.max_pitch_cmd == np.radians(10.0) and p.lookahead_time == 2.0 and p.min_speed == 12.0
    # the rate / angle loops come from the legacy ControllerConfig, not from the YAML's rate_control section (as in the reference)
    assert p.rate[0].kp == f32(1.3) and p.max_roll_rate == np.radians(180.0)
    # defaults + the repository's cascaded_pid.yaml == acb200_default_controller_params, byte for byte
    d = acb200.default_controller_params()
    assert bytes(cl.controller_params()) == bytes(d)
    assert bytes(cl.controller_params(ControllerConfig(), cl.load_controller_config("cascaded_pid.yaml"))) == bytes(d)
    m = cl.load_mission_config("square_pattern.yaml")
    assert (m.pattern_size, m.altitude, m.speed, m.guidance, m.dt) == (300.0, 100.0, 15.0, "PP", 0.01)
    assert cl.square_mission_waypoints(m)[2] == [300.0, 300.0, -100.0, 15.0]


@pytest.mark.skipif(not os.path.isdir("/root/reference/configs"), reason="reference checkout not present")
def test_loader_reads_the_reference_files_like_the_reference_loader():
    import dataclasses
    import importlib.util
    spec = importlib.util.spec_from_file_location("ref_config_loader", "/root/reference/controllers/config_loader.py")
    ref = importlib.util.module_from_spec(spec); spec.loader.exec_module(ref)
    for name in ("cascaded_pid.yaml", "default_gains.yaml", "jsbsim_gains.yaml"):
        ours = cl.load_controller_config(f"/root/reference/configs/controllers/{name}")
        assert dataclasses.asdict(ours) == dataclasses.asdict(ref.load_controller_config(name)), name
    ours = cl.load_mission_config("/root/reference/configs/missions/square_pattern.yaml")
    assert dataclasses.asdict(ours) == dataclasses.asdict(ref.load_mission_config("square_pattern.yaml"))
    # the repository's own copy of the C5 gains says what the reference's file says
    assert dataclasses.asdict(cl.load_controller_config("cascaded_pid.yaml")) == dataclasses.asdict(ref.load_controller_config("cascaded_pid.yaml"))
