"""The reference-shaped object API: the pybind-module mirror (pid_bindings), the agents (adapters) and the N = 1
backend.  CPU tests pin the host side against the reference's own known answers (tests/test_pid_bindings.py in the
reference: analytic cases) and against golden vectors generated from the compiled reference; the GPU tests fly
examples/01_hello_controls.py's loop (config C1) through the drop-in objects."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, rel_err


# ---- pid_bindings: the reference's tests/test_pid_bindings.py cases, same numbers --------------------------------
def _acb():
    from acb200 import pid_bindings as acb
    return acb


def test_bindings_struct_defaults_and_version():
    acb = _acb()
    g = acb.PIDGains(); assert (g.kp, g.ki, g.kd) == (0.0, 0.0, 0.0)
    g = acb.PIDGains(1.0, 0.5, 0.1); assert g.kp == 1.0 and g.ki == 0.5 and abs(g.kd - 0.1) < 1e-6
    c = acb.PIDConfig()
    assert (c.output_min, c.output_max, c.integral_min, c.integral_max) == (-1.0, 1.0, -10.0, 10.0)
    assert 0.0 <= c.derivative_filter_alpha <= 1.0 and c.derivative_filter_alpha == np.float32(0.1)
    assert acb.__version__ == "1.1.0"
    assert "kp=" in repr(g) and "MultiAxisPIDController(" in repr(acb.MultiAxisPIDController())


def test_bindings_analytic_cases():
    acb = _acb()
    cfg = acb.PIDConfig(); cfg.gains = acb.PIDGains(2.0, 0.0, 0.0)
    pid = acb.PIDController(cfg)
    assert pid.compute(setpoint=10.0, measurement=5.0, dt=0.01) == 1.0          # P-only saturates
    cfg = acb.PIDConfig(); cfg.gains = acb.PIDGains(0.0, 1.0, 0.0)
    pid = acb.PIDController(cfg)
    for _ in range(10):
        out = pid.compute(1.0, 0.0, 0.1)
    assert abs(out - 1.0) < 1e-5 and abs(pid.get_integral() - 1.0) < 1e-5       # I accumulates to 1.0
    pid.reset()
    assert (pid.get_error(), pid.get_integral(), pid.get_derivative(), pid.get_output()) == (0.0, 0.0, 0.0, 0.0)
    cfg = acb.PIDConfig(); cfg.gains = acb.PIDGains(10.0, 0.0, 0.0); cfg.output_min, cfg.output_max = -0.5, 0.5
    assert acb.PIDController(cfg).compute(10.0, 0.0, 0.01) == 0.5 and acb.PIDController(cfg).compute(-10.0, 0.0, 0.01) == -0.5
    cfg = acb.PIDConfig(); cfg.gains = acb.PIDGains(0.0, 1.0, 0.0); cfg.integral_min, cfg.integral_max = -5.0, 5.0
    cfg.output_min, cfg.output_max = -100.0, 100.0
    pid = acb.PIDController(cfg)
    for _ in range(100):
        pid.compute(10.0, 0.0, 0.1)
    assert pid.get_integral() == 5.0                                              # anti-windup clamp
    cfg = acb.PIDConfig(); cfg.gains = acb.PIDGains(0.0, 0.0, 1.0); cfg.derivative_filter_alpha = 1.0
    cfg.output_min, cfg.output_max = -100.0, 100.0
    pid = acb.PIDController(cfg)
    assert pid.compute(0.0, 0.0, 0.1) == 0.0 and abs(pid.compute(5.0, 0.0, 0.1) - 50.0) < 1e-4   # derivative 0 then 50
    # the constructor copies the config (C++ value semantics)
    cfg.gains = acb.PIDGains(9.0, 0.0, 0.0)
    assert pid.get_gains().kp == 0.0
    m = acb.MultiAxisPIDController(cfg, cfg, cfg)
    m.set_gains(7, acb.PIDGains(1, 1, 1)); assert m.get_gains(7).kp == 0.0       # invalid axis: ignored / default


def test_bindings_match_reference_known_answers_bit_exact():
    acb = _acb()
    g = np.load(os.path.join(GOLDEN, "pid_kat.npz"))
    for ci in range(g["cfg"].shape[0]):
        kp, ki, kd, omin, omax, imin, imax, alpha = [float(x) for x in g["cfg"][ci]]
        cfg = acb.PIDConfig(); cfg.gains = acb.PIDGains(kp, ki, kd)
        cfg.output_min, cfg.output_max, cfg.integral_min, cfg.integral_max, cfg.derivative_filter_alpha = omin, omax, imin, imax, alpha
        pid = acb.PIDController(cfg)
        for t in range(g["sp"].shape[1]):
            if np.isnan(g["sp"][ci, t]):
                pid.reset(); continue
            o = pid.compute(float(g["sp"][ci, t]), float(g["meas"][ci, t]), float(g["dt"][ci, t]))
            got = np.array([o, pid.get_error(), pid.get_integral(), pid.get_derivative()], np.float32)
            assert np.array_equal(got, g["out"][ci, t], equal_nan=True), (ci, t)


def test_agent_error_conventions():
    from acb200 import adapters as ad
    st = ad.AircraftState()
    with pytest.raises(AssertionError):
        ad.RateAgent().compute_action(ad.ControlCommand(mode=ad.ControlMode.ATTITUDE), st)
    with pytest.raises(ValueError):
        ad.RateAgent().compute_action(ad.ControlCommand(mode=ad.ControlMode.RATE, roll_rate=0.1), st)
    with pytest.raises(ValueError):
        ad.AttitudeAgent().compute_action(ad.ControlCommand(mode=ad.ControlMode.ATTITUDE, roll_angle=0.1), st)
    assert ad.wrap_angle(np.pi) == -np.pi and ad.wrap_angle(-np.pi) == -np.pi          # [-pi, pi)
    s = ad.SurfaceAgent().compute_action(ad.ControlCommand(mode=ad.ControlMode.SURFACE, elevator=2, aileron=-2, rudder=0.3, throttle=1.5))
    assert (s.elevator, s.aileron, s.rudder, s.throttle) == (1.0, -1.0, 0.3, 1.0)
    assert list(s.to_array()) == [1.0, -1.0, 0.3, 1.0]                                   # elevator, aileron, rudder, throttle


# ---- GPU: config C1 through the drop-in objects ---------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("name", ["rate_agent", "attitude_agent"])
def test_hello_controls_loop_matches_reference(name):
    """examples/01_hello_controls.py / tests/test_control_integration.py:34-74: agent.compute_action -> backend.set_controls
    -> backend.step, one aircraft, objects all the way; compared with the reference's trajectory (golden, aircraft 0)."""
    from acb200 import adapters as ad
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    level, hold, every = int(g["level"]), int(g["hold"]), int(g["every"])
    dt, dt_ctrl = float(g["dt"]), float(g["dt_ctrl"])
    k = 0
    s0 = g["s0"][:, k]
    backend = ad.SimulationAircraftBackendB200({"aircraft_type": "rc_plane"})
    agent = ad.RateAgent() if level == 4 else ad.AttitudeAgent()
    state = backend.reset(ad.AircraftState(time=s0[12], position=s0[0:3], velocity=s0[3:6], attitude=s0[6:9], angular_rate=s0[9:12]))
    steps = min(int(g["steps"]), 400)
    worst = 0.0
    for t in range(steps):
        c = g["cmd"][t // hold][:, k]
        if level == 4:
            cmd = ad.ControlCommand(mode=ad.ControlMode.RATE, roll_rate=c[0], pitch_rate=c[1], yaw_rate=c[2], throttle=c[3])
        else:
            yaw = None if np.isnan(c[2]) else (state.yaw if np.isinf(c[2]) else c[2])
            cmd = ad.ControlCommand(mode=ad.ControlMode.ATTITUDE, roll_angle=c[0], pitch_angle=c[1], yaw_angle=yaw, throttle=c[3])
        surfaces = agent.compute_action(cmd, state, dt=dt_ctrl)
        backend.set_controls(surfaces)
        state = backend.step(dt)
        if (t + 1) % every == 0:
            j = (t + 1) // every - 1
            got = np.concatenate([state.position, state.velocity, state.attitude, state.angular_rate, [state.time]])
            worst = max(worst, rel_err(got, g["states"][j][:, k]), rel_err(surfaces.to_array(), g["surfaces"][j][:, k]))
            assert worst <= 1e-9, (t + 1, worst)
    print(f"{name} through the N=1 objects: worst rel err {worst:.3e} over {steps} control steps")


@pytest.mark.gpu
def test_backend_contract():
    """The AircraftInterface facts the reference's tests/test_simulation.py pins (:46-62, 98-111, 229-375)."""
    from acb200 import adapters as ad
    b = ad.SimulationAircraftBackendB200({"aircraft_type": "rc_plane"})
    s = b.reset()
    assert s.altitude == 100.0 and s.airspeed == 20.0 and s.time == 0.0
    b.set_controls(ad.ControlSurfaces(throttle=0.5))
    s = b.step(0.01)
    assert abs(s.time - 0.01) < 1e-12
    s = b.step(0.1)
    assert abs(s.time - 0.11) < 1e-9                                  # sub-stepping: 100 substeps of 1 ms
    assert b.get_backend_type() == "simulation" and b.get_dt_nominal() == 0.01
    assert not b.is_real_hardware() and b.supports_reset()
    info = b.get_info()
    assert {"backend_type", "dt_nominal", "physics_engine", "dt_physics", "aircraft_mass", "max_thrust"} <= set(info)
    assert info["aircraft_mass"] == 8.0
    assert ad.SimulationAircraftBackendB200({"aircraft_type": "cessna"}).get_info()["aircraft_mass"] == 15.0
    assert ad.SimulationAircraftBackendB200({"dt_physics": 0.005})._dt_physics == 0.005
    with pytest.raises(ValueError):
        b.step(0.0)                                                   # simplified_6dof.py:241-245
    with pytest.raises(ValueError):
        ad.SimulationAircraftBackendB200({"dt_physics": 2.0}).step(2.0)   # one substep of 2 s > max_timestep
    # finite over 1000 steps with full-deflection controls (test_simulation.py:187-209)
    b.reset()
    b.set_controls(ad.ControlSurfaces(elevator=1.0, aileron=1.0, rudder=1.0, throttle=1.0))
    for _ in range(200):
        s = b.step(0.01)
    assert np.all(np.isfinite(s.position)) and np.all(np.isfinite(s.velocity)) and np.all(np.isfinite(s.attitude))
