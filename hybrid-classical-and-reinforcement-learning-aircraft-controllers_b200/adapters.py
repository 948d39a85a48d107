"""Single-aircraft adapters with the reference's own object API on top of the batched CUDA path (N = 1).

``SimulationAircraftBackendB200`` keeps the ``AircraftInterface`` contract of ``SimulationAircraftBackend``
(interfaces/aircraft.py:19-171, simulation/simulation_backend.py:13-170): ``reset / set_controls / step /
get_state / close / get_backend_type / get_dt_nominal / is_real_hardware / supports_reset / get_info`` with
dataclass values in and out.  ``RateAgent`` / ``AttitudeAgent`` / ``SurfaceAgent`` keep ``compute_action(command,
state, dt)`` (controllers/rate_agent.py:65-124, attitude_agent.py:86-154, surface_agent.py:50-96) and run on the
library's fp32 PID core through ``pid_bindings`` -- the same arithmetic the fused kernels inline.

Objects are duck-typed: the reference's own ``AircraftState`` / ``ControlSurfaces`` / ``ControlCommand`` instances
are accepted wherever the ones defined here are.  For more than a handful of aircraft use ``AircraftFleet`` /
``BatchedRateControlEnv``: one object per aircraft is exactly the overhead the batched path removes.
"""
from __future__ import annotations

import enum
import math
from dataclasses import dataclass, field
from typing import Optional

import numpy as np

from . import pid_bindings as acb


class ControlMode(enum.IntEnum):           # controllers/types.py:9-24
    WAYPOINT = 1
    HSA = 2
    ATTITUDE = 3
    RATE = 4
    SURFACE = 5

    # equal to the reference's own (non-int) ``ControlMode`` members of the same value, so that either side's objects can be
    # compared with ``==`` (reference tests assert ``agent.get_control_level() == ControlMode.RATE`` with THEIR enum)
    def __eq__(self, other):
        v = getattr(other, "value", other)
        return int(self) == v if isinstance(v, int) else NotImplemented

    def __ne__(self, other):
        r = self.__eq__(other)
        return r if r is NotImplemented else not r

    __hash__ = int.__hash__


@dataclass
class AircraftState:                       # controllers/types.py:27-96
    time: float = 0.0
    position: np.ndarray = field(default_factory=lambda: np.zeros(3))      # N, E, D
    velocity: np.ndarray = field(default_factory=lambda: np.zeros(3))      # u, v, w (body)
    attitude: np.ndarray = field(default_factory=lambda: np.zeros(3))      # roll, pitch, yaw
    angular_rate: np.ndarray = field(default_factory=lambda: np.zeros(3))  # p, q, r
    airspeed: float = 0.0
    altitude: float = 0.0
    ground_speed: float = 0.0
    heading: float = 0.0

    roll = property(lambda s: s.attitude[0])
    pitch = property(lambda s: s.attitude[1])
    yaw = property(lambda s: s.attitude[2])
    p = property(lambda s: s.angular_rate[0])
    q = property(lambda s: s.angular_rate[1])
    r = property(lambda s: s.angular_rate[2])
    north = property(lambda s: s.position[0])
    east = property(lambda s: s.position[1])
    down = property(lambda s: s.position[2])


@dataclass
class ControlSurfaces:                     # controllers/types.py:173-202
    elevator: float = 0.0
    aileron: float = 0.0
    rudder: float = 0.0
    throttle: float = 0.0

    def to_array(self) -> np.ndarray:
        return np.array([self.elevator, self.aileron, self.rudder, self.throttle])

    @classmethod
    def from_array(cls, arr):
        return cls(elevator=arr[0], aileron=arr[1], rudder=arr[2], throttle=arr[3])


@dataclass
class ControlCommand:                      # controllers/types.py:135-170
    mode: ControlMode
    timestamp: float = 0.0
    waypoint: Optional[object] = None
    heading: Optional[float] = None
    speed: Optional[float] = None
    altitude: Optional[float] = None
    roll_angle: Optional[float] = None
    pitch_angle: Optional[float] = None
    yaw_angle: Optional[float] = None
    roll_rate: Optional[float] = None
    pitch_rate: Optional[float] = None
    yaw_rate: Optional[float] = None
    throttle: Optional[float] = None
    elevator: Optional[float] = None
    aileron: Optional[float] = None
    rudder: Optional[float] = None


@dataclass
class PIDGains:                            # controllers/types.py:277-291
    kp: float = 0.0
    ki: float = 0.0
    kd: float = 0.0
    i_limit: float = 25.0


@dataclass
class Waypoint:                            # controllers/types.py:99-132
    north: float
    east: float
    down: float
    speed: Optional[float] = None
    heading: Optional[float] = None

    @property
    def altitude(self) -> float:
        return -self.down

    @classmethod
    def from_ned(cls, north, east, down, **kw):
        return cls(north=north, east=east, down=down, **kw)

    @classmethod
    def from_altitude(cls, north, east, altitude, **kw):
        return cls(north=north, east=east, down=-altitude, **kw)


@dataclass
class ControllerConfig:                    # controllers/types.py:294-350
    roll_angle_gains: PIDGains = field(default_factory=lambda: PIDGains(kp=8.0, ki=2.0, kd=0.3))
    pitch_angle_gains: PIDGains = field(default_factory=lambda: PIDGains(kp=6.0, ki=1.5, kd=0.2))
    roll_rate_gains: PIDGains = field(default_factory=lambda: PIDGains(kp=1.3, ki=0.4, kd=0.012))
    pitch_rate_gains: PIDGains = field(default_factory=lambda: PIDGains(kp=0.6, ki=0.2, kd=0.008))
    yaw_gains: PIDGains = field(default_factory=lambda: PIDGains(kp=1.6, ki=0.15, kd=0.01))
    heading_gains: PIDGains = field(default_factory=lambda: PIDGains(kp=1.0, ki=0.05, kd=0.2))      # only i_limit is read (hsa_agent.py:64-65)
    speed_gains: PIDGains = field(default_factory=lambda: PIDGains(kp=0.1, ki=0.01, kd=0.0))
    altitude_gains: PIDGains = field(default_factory=lambda: PIDGains(kp=0.2, ki=0.01, kd=0.1))
    max_roll: float = 30.0          # degrees
    max_pitch: float = 30.0
    max_roll_rate: float = 180.0    # degrees / s
    max_pitch_rate: float = 180.0
    max_yaw_rate: float = 160.0
    rate_loop_dt: float = 0.001
    dt: float = 0.01
    max_bank_angle_hsa: float = 25.0
    baseline_throttle: float = 0.1
    waypoint_acceptance_radius: float = 40.0
    proportional_navigation_gain: float = 3.0


def _same_mode(a, b) -> bool:
    return int(getattr(a, "value", a)) == int(getattr(b, "value", b))


def create_pid_config(gains, output_min: float = -1.0, output_max: float = 1.0) -> acb.PIDConfig:
    """controllers/utils/pid_utils.py:16-46"""
    c = acb.PIDConfig()
    c.gains = acb.PIDGains(gains.kp, gains.ki, gains.kd)
    c.integral_min, c.integral_max = -gains.i_limit, gains.i_limit
    c.output_min, c.output_max = output_min, output_max
    return c


def wrap_angle(a: float) -> float:
    """controllers/utils/pid_utils.py:49-65: floored modulo, range [-pi, pi)"""
    return (a + np.pi) % (2 * np.pi) - np.pi


def validate_command(command, mode: str, required: list) -> None:
    """controllers/utils/pid_utils.py:68-95"""
    for f in required:
        if getattr(command, f, None) is None:
            raise ValueError(f"Invalid {mode} command: '{f}' is None. All of {required} must be specified.")


class SurfaceAgent:
    """controllers/surface_agent.py:50-96: clip only."""

    def __init__(self, config=None):
        self.config = config

    def get_control_level(self):
        return ControlMode.SURFACE

    def compute_action(self, command, state=None, dt=None) -> ControlSurfaces:
        assert _same_mode(command.mode, ControlMode.SURFACE), f"Surface agent expects SURFACE mode, got {command.mode}"
        validate_command(command, "SURFACE", ["elevator", "aileron", "rudder", "throttle"])
        return ControlSurfaces(elevator=float(np.clip(command.elevator, -1.0, 1.0)), aileron=float(np.clip(command.aileron, -1.0, 1.0)),
                               rudder=float(np.clip(command.rudder, -1.0, 1.0)), throttle=float(np.clip(command.throttle, 0.0, 1.0)))

    def reset(self):
        pass


class RateAgent:
    """controllers/rate_agent.py:32-128"""

    def __init__(self, config=None):
        self.config = config or ControllerConfig()
        c = self.config
        self.rate_controller = acb.MultiAxisPIDController(create_pid_config(c.roll_rate_gains), create_pid_config(c.pitch_rate_gains),
                                                          create_pid_config(c.yaw_gains))
        self.max_roll_rate = np.radians(c.max_roll_rate)
        self.max_pitch_rate = np.radians(c.max_pitch_rate)
        self.max_yaw_rate = np.radians(c.max_yaw_rate)

    def get_control_level(self):
        return ControlMode.RATE

    def compute_action(self, command, state, dt: float = None) -> ControlSurfaces:
        assert _same_mode(command.mode, ControlMode.RATE), f"Rate agent expects RATE mode, got {command.mode}"
        validate_command(command, "RATE", ["roll_rate", "pitch_rate", "yaw_rate"])
        sp = acb.Vector3(np.clip(command.roll_rate, -self.max_roll_rate, self.max_roll_rate),
                         np.clip(command.pitch_rate, -self.max_pitch_rate, self.max_pitch_rate),
                         np.clip(command.yaw_rate, -self.max_yaw_rate, self.max_yaw_rate))
        meas = acb.Vector3(state.p, state.q, state.r)
        out = self.rate_controller.compute(sp, meas, dt if dt is not None else self.config.rate_loop_dt)
        return ControlSurfaces(aileron=float(np.clip(out.roll, -1.0, 1.0)), elevator=float(np.clip(-out.pitch, -1.0, 1.0)),
                               rudder=float(np.clip(-out.yaw, -1.0, 1.0)), throttle=float(np.clip(command.throttle, 0.0, 1.0)))

    def reset(self):
        self.rate_controller.reset()


class AttitudeAgent:
    """controllers/attitude_agent.py:33-159"""

    def __init__(self, config=None):
        self.config = config or ControllerConfig()
        c = self.config
        lim = lambda deg: (-np.radians(deg), np.radians(deg))
        self.angle_controller = acb.MultiAxisPIDController(create_pid_config(c.roll_angle_gains, *lim(c.max_roll_rate)),
                                                           create_pid_config(c.pitch_angle_gains, *lim(c.max_pitch_rate)),
                                                           create_pid_config(c.yaw_gains, *lim(c.max_yaw_rate)))
        self.rate_agent = RateAgent(c)
        self.max_roll, self.max_pitch = np.radians(c.max_roll), np.radians(c.max_pitch)
        self._rates = (np.radians(c.max_roll_rate), np.radians(c.max_pitch_rate), np.radians(c.max_yaw_rate))

    def get_control_level(self):
        return ControlMode.ATTITUDE

    def compute_action(self, command, state, dt: float = 0.01) -> ControlSurfaces:
        assert _same_mode(command.mode, ControlMode.ATTITUDE), f"Attitude agent expects ATTITUDE mode, got {command.mode}"
        validate_command(command, "ATTITUDE", ["roll_angle", "pitch_angle"])
        roll_cmd = np.clip(command.roll_angle, -self.max_roll, self.max_roll)
        pitch_cmd = np.clip(command.pitch_angle, -self.max_pitch, self.max_pitch)
        yaw_cmd = wrap_angle(command.yaw_angle) if command.yaw_angle is not None else 0.0
        out = self.angle_controller.compute(acb.Vector3(roll_cmd, pitch_cmd, yaw_cmd),
                                            acb.Vector3(state.roll, state.pitch, wrap_angle(state.yaw)), dt)
        rate_cmd = ControlCommand(mode=ControlMode.RATE, roll_rate=np.clip(out.roll, -self._rates[0], self._rates[0]),
                                  pitch_rate=np.clip(out.pitch, -self._rates[1], self._rates[1]),
                                  yaw_rate=np.clip(out.yaw, -self._rates[2], self._rates[2]), throttle=command.throttle)
        return self.rate_agent.compute_action(rate_cmd, state, dt=dt)

    def reset(self):
        self.angle_controller.reset()
        self.rate_agent.reset()


class SimulationAircraftBackendB200:
    """``SimulationAircraftBackend`` (simulation/simulation_backend.py:13-170) over a one-aircraft ``AircraftFleet``.

    precision 64 (default) is the parity build: trajectories match the reference within 1e-9.
    config keys: ``aircraft_type`` ('rc_plane' | 'cessna'), ``dt_physics`` (default 0.001), ``params`` (an object with
    the reference's AircraftParams attribute names, or an ``_cabi.AircraftParamsC``).
    """

    def __init__(self, config: Optional[dict] = None, precision: int = 64, device=0):
        from . import _cabi
        from .fleet import AircraftFleet
        import torch
        self._torch = torch
        config = dict(config or {})
        params = config.get("params")
        if params is not None and not isinstance(params, _cabi.AircraftParamsC):
            p = _cabi.default_aircraft_params()
            for f in _cabi.AIRCRAFT_FIELDS:          # AircraftParams dataclass -> POD (simplified_6dof.py:31-117)
                if hasattr(params, f):
                    setattr(p, f, float(getattr(params, f)))
            config["params"] = p
        self._fleet = AircraftFleet(1, precision=precision, device=device, config=config)   # raises ValueError like __post_init__
        self._dt_physics = self._fleet.dt_physics
        self._state = None

    # -- AircraftInterface -----------------------------------------------------------------------
    def _read_state(self) -> AircraftState:
        s = self._fleet.state[:, 0].double().cpu().numpy()
        d = self._fleet.derived[:, 0].double().cpu().numpy()
        return AircraftState(time=float(s[12]), position=s[0:3].copy(), velocity=s[3:6].copy(), attitude=s[6:9].copy(),
                             angular_rate=s[9:12].copy(), airspeed=float(d[0]), altitude=float(d[1]),
                             ground_speed=float(d[2]), heading=float(d[3]))

    def reset(self, initial_state=None) -> AircraftState:
        if initial_state is None:
            self._fleet.reset()
        else:
            s = np.concatenate([np.asarray(initial_state.position, float), np.asarray(initial_state.velocity, float),
                                np.asarray(initial_state.attitude, float), np.asarray(initial_state.angular_rate, float),
                                [float(initial_state.time)]])
            self._fleet.reset(self._torch.as_tensor(s.reshape(13, 1)))
        self._state = self._read_state()
        return self._state

    def set_controls(self, surfaces) -> None:
        c = [surfaces.elevator, surfaces.aileron, surfaces.rudder, surfaces.throttle]
        self._fleet.set_controls(self._torch.as_tensor(np.asarray(c, float).reshape(4, 1)))

    def step(self, dt: float) -> AircraftState:
        self._fleet.step(dt)                       # ValueError for a bad dt (simplified_6dof.py:241-245)
        self._state = self._read_state()
        return self._state

    def get_state(self) -> AircraftState:
        if self._state is None:
            self._state = self._read_state()
        return self._state

    def close(self) -> None:
        pass

    def get_backend_type(self) -> str:
        return "simulation"

    def get_dt_nominal(self) -> float:
        return 0.01

    def is_real_hardware(self) -> bool:
        return False

    def supports_reset(self) -> bool:
        return True

    def get_info(self) -> dict:
        p = self._fleet.aircraft_params
        return {"backend_type": "simulation", "dt_nominal": 0.01, "physics_engine": "simplified_6dof",
                "dt_physics": self._dt_physics, "aircraft_mass": p.mass, "max_thrust": p.max_thrust}

    def __repr__(self) -> str:
        return f"SimulationAircraftBackendB200(physics=Simplified6DOF/CUDA, dt={self._dt_physics})"
