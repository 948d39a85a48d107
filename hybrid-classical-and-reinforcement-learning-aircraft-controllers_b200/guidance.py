"""Levels 2 and 1 of the cascade as drop-in objects: ``HSAAgent``, ``WaypointAgent``, ``MissionPlanner``.

One-aircraft mirrors of controllers/hsa_agent.py:17-238, controllers/waypoint_agent.py:13-272 and
controllers/mission_planner.py:18-300 with the reference's constructor signatures, ``BaseAgent`` methods
(``get_control_level / compute_action(command, state, dt) / reset``), attributes the reference's tests read
(``.heading_pid .energy_pid .balance_pid .attitude_agent``; ``.hsa_agent .guidance_type .acceptance_radius
.last_heading_cmd``; ``.current_waypoint_index .waypoints_reached .state``) and exception types.  Like the rate /
attitude adapters they run the library's fp32 PID core (``pid_bindings`` -> ``acb200_pid_compute_f32``) -- the arithmetic
``step_guided_kernel`` inlines for a whole fleet (``AircraftFleet.step_hsa`` / ``set_mission`` + ``step_waypoint``);
tests/test_gpu_guidance_objects.py flies the two side by side.  They read ``state.heading / airspeed / altitude`` as
given, as the reference does (a caller may hand in a state whose derived fields it set itself).
"""
from __future__ import annotations

import enum
from typing import List, Optional

import numpy as np

from . import pid_bindings as acb
from .adapters import AttitudeAgent, ControlCommand, ControlMode, ControlSurfaces, ControllerConfig, Waypoint, _same_mode, \
    validate_command, wrap_angle
from .config_loader import FlightControlConfig, GuidanceConfig, HSAConfig, TECSConfig

__all__ = ["HSAAgent", "WaypointAgent", "MissionPlanner", "MissionState", "Waypoint"]


def _scalar_pid(gains, i_limit: float, out_limit: float) -> acb.PIDController:
    c = acb.PIDConfig()
    c.gains = acb.PIDGains(kp=gains.kp, ki=gains.ki, kd=gains.kd)
    c.integral_min, c.integral_max = -i_limit, i_limit
    c.output_min, c.output_max = -out_limit, out_limit
    return acb.PIDController(c)


def _wrap_pi(a):
    """the reference's ``arctan2(sin(a), cos(a))`` idiom (waypoint_agent.py:132,147,185)"""
    return np.arctan2(np.sin(a), np.cos(a))


class HSAAgent:
    """Level 2 (hsa_agent.py:17-238): heading PID -> bank; TECS energy PID -> throttle, balance PID -> pitch; level 3 inside."""

    def __init__(self, config=None, flight_config: Optional[FlightControlConfig] = None):
        self.config = config if config is not None else ControllerConfig()
        hsa_cfg = flight_config.hsa if flight_config is not None else HSAConfig()
        tecs_cfg = hsa_cfg.tecs if flight_config is not None else TECSConfig()
        self.max_bank_rad = np.radians(hsa_cfg.max_bank_angle)
        heading_i = getattr(getattr(self.config, "heading_gains", None), "i_limit", 25.0)
        self.heading_pid = _scalar_pid(hsa_cfg.heading_gains, heading_i, self.max_bank_rad)
        self.energy_pid = _scalar_pid(tecs_cfg.energy_gains, 10.0, 0.5)
        self.balance_pid = _scalar_pid(tecs_cfg.balance_gains, 5.0, np.radians(tecs_cfg.max_pitch_command))
        self.attitude_agent = AttitudeAgent(self.config)
        self.baseline_throttle = tecs_cfg.baseline_throttle
        self.load_factor_gain = tecs_cfg.load_factor_gain
        self.max_pitch_command = tecs_cfg.max_pitch_command
        self._max_pitch_cmd_rad = np.radians(10.0)                     # :108, a constant in the reference

    def get_control_level(self):
        return ControlMode.HSA

    def compute_action(self, command, state, dt: float = 0.01) -> ControlSurfaces:
        assert _same_mode(command.mode, ControlMode.HSA), f"HSA agent expects HSA mode, got {command.mode}"
        validate_command(command, "HSA", ["heading", "altitude", "speed"])
        # heading: the PID sees the wrapped error through a virtual setpoint (:146-153)
        heading_error = wrap_angle(command.heading - state.heading)
        roll_angle = self.heading_pid.compute(state.heading + heading_error, state.heading, dt)
        roll_angle = np.clip(roll_angle, -self.max_bank_rad, self.max_bank_rad)
        # TECS on specific energies (:160-206)
        g, h, V, h_cmd, V_cmd = 9.81, state.altitude, state.airspeed, command.altitude, command.speed
        e_total, e_total_cmd = g * h + 0.5 * V ** 2, g * h_cmd + 0.5 * V_cmd ** 2
        e_bal, e_bal_cmd = g * h - 0.5 * V ** 2, g * h_cmd - 0.5 * V_cmd ** 2
        throttle = np.clip(self.baseline_throttle + self.energy_pid.compute(e_total_cmd, e_total, dt), 0.0, 1.0)
        pitch_angle = self.balance_pid.compute(e_bal_cmd, e_bal, dt)
        cos_roll = np.cos(roll_angle)
        if abs(cos_roll) > 0.01:                                        # load-factor compensation in turns (:209-212)
            pitch_angle += self.load_factor_gain * (1.0 / cos_roll - 1.0)
        pitch_angle = np.clip(pitch_angle, -self._max_pitch_cmd_rad, self._max_pitch_cmd_rad)
        att = ControlCommand(mode=ControlMode.ATTITUDE, roll_angle=roll_angle, pitch_angle=pitch_angle, yaw_angle=state.yaw,
                             throttle=throttle)
        return self.attitude_agent.compute_action(att, state, dt)

    def reset(self):
        self.heading_pid.reset(); self.energy_pid.reset(); self.balance_pid.reset()
        self.attitude_agent.reset()

    def __repr__(self) -> str:
        return "HSAAgent(level=2, inner=AttitudeAgent)"


class WaypointAgent:
    """Level 1 (waypoint_agent.py:13-272): LOS or pure-pursuit guidance -> HSA command -> level 2 inside."""

    def __init__(self, config=None, guidance_type: str = "LOS", flight_config: Optional[FlightControlConfig] = None):
        self.config = config if config is not None else ControllerConfig()
        self.guidance_type = guidance_type
        self.guidance = flight_config.guidance if flight_config is not None else GuidanceConfig()
        self.max_bank_angle = flight_config.hsa.max_bank_angle if flight_config is not None else 25.0
        self.hsa_agent = HSAAgent(self.config, flight_config)
        self.acceptance_radius = self.guidance.acceptance_radius
        self.N = self.guidance.pn_gain
        self._error_buf = np.zeros(3)
        self.last_heading_cmd = None
        self.max_heading_rate = None                                     # rate limiting is off in the reference (:70)

    def get_control_level(self):
        return ControlMode.WAYPOINT

    def _compute_error(self, state, waypoint) -> np.ndarray:
        self._error_buf[0] = waypoint.north - state.north
        self._error_buf[1] = waypoint.east - state.east
        self._error_buf[2] = waypoint.down - state.down
        return self._error_buf

    def compute_action(self, command, state, dt: float = 0.01) -> ControlSurfaces:
        assert _same_mode(command.mode, ControlMode.WAYPOINT), f"Waypoint agent expects WAYPOINT mode, got {command.mode}"
        wp, G = command.waypoint, self.guidance
        error = self._compute_error(state, wp)
        dist = np.linalg.norm(error[:2])
        los = np.arctan2(error[1], error[0])
        heading_cmd = los
        if self.guidance_type == "LOS":
            # turn anticipation: bank early when close and the turn is wide (:121-144)
            V = max(state.airspeed, 10.0)
            turn_radius = V ** 2 / (9.81 * np.tan(np.radians(G.los_max_bank)))
            herr = _wrap_pi(heading_cmd - state.heading)
            anticipation = turn_radius * abs(herr) / np.radians(90)
            if dist < anticipation and abs(herr) > np.radians(20):
                blend = 1.0 - (dist / anticipation)
                heading_cmd = _wrap_pi(heading_cmd + blend * np.radians(G.los_lead_angle) * np.sign(herr))
        elif self.guidance_type in ("PP", "PURE_PURSUIT"):
            # carrot on the line to the waypoint at a speed-scaled, proximity-shortened, clamped lookahead (:146-178)
            V = max(state.airspeed, 10.0)
            turn_radius = V ** 2 / (9.81 * np.tan(np.radians(self.max_bank_angle)))
            lookahead = G.lookahead_time * V
            proximity = G.proximity_scale_distance * turn_radius
            if dist < proximity:
                lookahead *= 0.6 + 0.4 * (dist / proximity)
            lookahead = np.clip(lookahead, G.lookahead_min, G.lookahead_max)
            if dist > 1.0:
                carrot = (error[:2] / dist) * min(lookahead, dist)
                heading_cmd = np.arctan2(carrot[1], carrot[0])
        heading_cmd = _wrap_pi(heading_cmd)
        if self.max_heading_rate is not None and self.last_heading_cmd is not None:
            change = _wrap_pi(heading_cmd - self.last_heading_cmd)
            if abs(change) > self.max_heading_rate * dt:
                heading_cmd = _wrap_pi(self.last_heading_cmd + np.sign(change) * self.max_heading_rate * dt)
        self.last_heading_cmd = heading_cmd
        # speed: the waypoint's or the current one, reduced for sharp turns close to the waypoint (:207-228)
        speed_cmd = wp.speed if wp.speed is not None else state.airspeed
        herr = abs(los - state.heading)
        herr = abs(np.arctan2(np.sin(herr), np.cos(herr)))
        if dist < G.turn_threshold_distance and herr > np.radians(G.turn_threshold_angle):
            speed_cmd = speed_cmd * (1.0 - G.max_speed_reduction * (1.0 - dist / G.turn_threshold_distance))
            speed_cmd = max(speed_cmd, G.min_speed)
        hsa_cmd = ControlCommand(mode=ControlMode.HSA, heading=heading_cmd, speed=speed_cmd, altitude=wp.altitude)
        return self.hsa_agent.compute_action(hsa_cmd, state, dt)

    def reached_waypoint(self, state, waypoint) -> bool:
        return bool(np.linalg.norm(self._compute_error(state, waypoint)) < self.acceptance_radius)

    def reset(self):
        self.last_heading_cmd = None
        self.hsa_agent.reset()

    def __repr__(self) -> str:
        return f"WaypointAgent(level=1, guidance={self.guidance_type}, inner=HSAAgent)"


class MissionState(enum.Enum):                 # mission_planner.py:10-15
    IDLE = "idle"
    ACTIVE = "active"
    COMPLETE = "complete"
    ABORTED = "aborted"


class MissionPlanner:
    """Waypoint sequencing with auto-advance (mission_planner.py:18-300)."""

    def __init__(self, waypoints: List[Waypoint], acceptance_radius: float = 10.0, default_speed: Optional[float] = None):
        if not waypoints:
            raise ValueError("Mission must have at least one waypoint")
        self.waypoints = waypoints
        self.acceptance_radius = acceptance_radius
        self.default_speed = default_speed
        self.reset()

    def start(self):
        if self.state == MissionState.IDLE:
            self.state = MissionState.ACTIVE
            self.current_waypoint_index = 0
            self.waypoints_reached = 0

    def abort(self):
        self.state = MissionState.ABORTED

    def reset(self):
        self.state = MissionState.IDLE
        self.current_waypoint_index = 0
        self.waypoints_reached = 0
        self.waypoint_arrival_times = []
        self.waypoint_distances = []
        self.mission_start_time = None
        self.mission_end_time = None

    def get_current_waypoint(self) -> Optional[Waypoint]:
        if self.state != MissionState.ACTIVE or self.current_waypoint_index >= len(self.waypoints):
            return None
        return self.waypoints[self.current_waypoint_index]

    def get_waypoint_command(self) -> Optional[ControlCommand]:
        wp = self.get_current_waypoint()
        return None if wp is None else ControlCommand(mode=ControlMode.WAYPOINT, waypoint=wp)

    def _calculate_distance_to_waypoint(self, state, waypoint) -> float:
        return float(np.linalg.norm(np.array([waypoint.north - state.north, waypoint.east - state.east, waypoint.down - state.down])))

    def _is_waypoint_reached(self, state, waypoint) -> bool:
        return self._calculate_distance_to_waypoint(state, waypoint) < self.acceptance_radius

    def update(self, state) -> bool:
        """True when the current waypoint was reached (3-D distance below the radius) and the index advanced (:128-171)."""
        if self.state != MissionState.ACTIVE:
            return False
        if self.mission_start_time is None:
            self.mission_start_time = state.time
        wp = self.get_current_waypoint()
        if wp is None or not self._is_waypoint_reached(state, wp):
            return False
        self.waypoint_arrival_times.append(state.time)
        self.waypoints_reached += 1
        self.waypoint_distances.append(self._calculate_distance_to_waypoint(state, wp))
        self.current_waypoint_index += 1
        if self.current_waypoint_index >= len(self.waypoints):
            self.state = MissionState.COMPLETE
            self.mission_end_time = state.time
        return True

    def get_distance_to_current_waypoint(self, state) -> Optional[float]:
        wp = self.get_current_waypoint()
        return None if wp is None else self._calculate_distance_to_waypoint(state, wp)

    def get_progress_percentage(self) -> float:
        return 100.0 if not self.waypoints else (self.waypoints_reached / len(self.waypoints)) * 100.0

    def get_total_mission_distance(self) -> float:
        total = 0.0
        for a, b in zip(self.waypoints, self.waypoints[1:]):
            total += float(np.linalg.norm([b.north - a.north, b.east - a.east, b.down - a.down]))
        return total

    def get_mission_duration(self) -> Optional[float]:
        if self.mission_start_time is None or self.mission_end_time is None:
            return None
        return self.mission_end_time - self.mission_start_time

    def is_active(self) -> bool:
        return self.state == MissionState.ACTIVE

    def is_complete(self) -> bool:
        return self.state == MissionState.COMPLETE

    def is_aborted(self) -> bool:
        return self.state == MissionState.ABORTED

    def get_summary(self) -> dict:
        return {"state": self.state.value, "total_waypoints": len(self.waypoints), "waypoints_reached": self.waypoints_reached,
                "progress_percent": self.get_progress_percentage(), "current_waypoint_index": self.current_waypoint_index,
                "total_distance_m": self.get_total_mission_distance(), "mission_duration_s": self.get_mission_duration(),
                "acceptance_radius_m": self.acceptance_radius, "waypoint_arrival_times": self.waypoint_arrival_times,
                "waypoint_distances": self.waypoint_distances}

    def __repr__(self) -> str:
        return (f"MissionPlanner(state={self.state.value}, waypoint={self.current_waypoint_index}/{len(self.waypoints)}, "
                f"reached={self.waypoints_reached})")

    # ---- the same mission for a whole fleet -------------------------------------------------------------------------
    def waypoint_table(self) -> np.ndarray:
        """[n_wp][4] = north, east, down, speed (NaN = None): what ``AircraftFleet.set_mission`` / ``acb200_mission`` take."""
        return np.array([[w.north, w.east, w.down, np.nan if w.speed is None else w.speed] for w in self.waypoints], dtype=np.float64)
