"""YAML -> controller / mission configuration, and configuration -> the C ABI's ``acb200_controller_params``.

Host-side mirror of the reference's ``controllers/config_loader.py`` (dataclasses :28-109, ``load_controller_config``
:112-207, ``load_mission_config`` :224-260): same class names, field names, YAML keys and defaults, so that
``configs/controllers/cascaded_pid.yaml`` and ``configs/missions/*.yaml`` of a reference checkout load unchanged (pass
the path, or a bare file name that is looked up under ``config_dir`` / ``$ACB200_CONFIG_DIR`` / this repository's
``configs/``).  ``controller_params`` is the one new function: it folds a legacy ``ControllerConfig`` and an optional
``FlightControlConfig`` into the POD the kernels read, exactly as the reference agents build their PIDs
(rate_agent.py:41-55, attitude_agent.py:43-71, hsa_agent.py:47-108, waypoint_agent.py:49-63).
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass, field
from pathlib import Path
from typing import Any, Dict, Optional

import yaml

from . import _cabi

_REPO_CONFIGS = Path(__file__).resolve().parent.parent / "configs"


def _config_dir(config_dir=None) -> Path:
    return Path(config_dir or os.environ.get("ACB200_CONFIG_DIR") or _REPO_CONFIGS)


def load_yaml(filepath) -> Dict[str, Any]:
    with open(filepath, "r") as f:
        return yaml.safe_load(f) or {}


@dataclass
class PIDGains:                                    # config_loader.py:28-37 (no i_limit here: that lives in types.PIDGains)
    kp: float = 0.0
    ki: float = 0.0
    kd: float = 0.0

    @classmethod
    def from_dict(cls, d: Dict[str, float]) -> "PIDGains":
        return cls(kp=d.get("kp", 0.0), ki=d.get("ki", 0.0), kd=d.get("kd", 0.0))


@dataclass
class GuidanceConfig:                              # config_loader.py:40-64
    acceptance_radius: float = 40.0
    lookahead_time: float = 1.2
    lookahead_min: float = 12.0
    lookahead_max: float = 30.0
    proximity_scale_distance: float = 2.0
    los_max_bank: float = 20.0
    los_lead_angle: float = 30.0
    turn_threshold_angle: float = 60.0
    turn_threshold_distance: float = 60.0
    max_speed_reduction: float = 0.15
    min_speed: float = 14.0
    pn_gain: float = 3.0


@dataclass
class TECSConfig:                                  # config_loader.py:66-74
    energy_gains: PIDGains = field(default_factory=lambda: PIDGains(0.12, 0.03, 0.03))
    balance_gains: PIDGains = field(default_factory=lambda: PIDGains(0.06, 0.005, 0.03))
    max_pitch_command: float = 15.0
    baseline_throttle: float = 0.1
    load_factor_gain: float = 0.05


@dataclass
class HSAConfig:                                   # config_loader.py:76-83
    heading_gains: PIDGains = field(default_factory=lambda: PIDGains(1.0, 0.05, 0.2))
    max_bank_angle: float = 25.0
    tecs: TECSConfig = field(default_factory=TECSConfig)


@dataclass
class FlightControlConfig:                         # config_loader.py:85-109
    outer_loop_dt: float = 0.01
    inner_loop_dt: float = 0.001
    roll_rate_gains: PIDGains = field(default_factory=lambda: PIDGains(1.3, 0.4, 0.012))
    pitch_rate_gains: PIDGains = field(default_factory=lambda: PIDGains(0.6, 0.2, 0.008))
    yaw_rate_gains: PIDGains = field(default_factory=lambda: PIDGains(1.6, 0.15, 0.01))
    max_roll_rate: float = 200.0
    max_pitch_rate: float = 100.0
    max_yaw_rate: float = 60.0
    roll_angle_gains: PIDGains = field(default_factory=lambda: PIDGains(8.0, 2.0, 0.3))
    pitch_angle_gains: PIDGains = field(default_factory=lambda: PIDGains(6.0, 1.5, 0.2))
    max_roll: float = 30.0
    max_pitch: float = 20.0
    hsa: HSAConfig = field(default_factory=HSAConfig)
    guidance: GuidanceConfig = field(default_factory=GuidanceConfig)


def _resolve(config_file, sub: str, config_dir=None) -> Path:
    p = Path(config_file)
    return p if (p.is_absolute() or p.exists()) else _config_dir(config_dir) / sub / config_file


def load_controller_config(config_file: str = "cascaded_pid.yaml", config_dir=None) -> FlightControlConfig:
    """config_loader.py:112-207: every key optional, the dataclass defaults otherwise; a missing file gives the defaults
    with a warning, as in the reference."""
    filepath = _resolve(config_file, "controllers", config_dir)
    config = FlightControlConfig()
    if not filepath.exists():
        print(f"Warning: Config file {filepath} not found, using defaults")
        return config
    data = load_yaml(filepath)
    t = data.get("timing") or {}
    config.outer_loop_dt = t.get("outer_loop_dt", 0.01)
    config.inner_loop_dt = t.get("inner_loop_dt", 0.001)
    rc = data.get("rate_control") or {}
    for key, attr in (("roll", "roll_rate_gains"), ("pitch", "pitch_rate_gains"), ("yaw", "yaw_rate_gains")):
        if key in rc:
            setattr(config, attr, PIDGains.from_dict(rc[key]))
    if "limits" in rc:
        config.max_roll_rate = rc["limits"].get("max_roll_rate", 200.0)
        config.max_pitch_rate = rc["limits"].get("max_pitch_rate", 100.0)
        config.max_yaw_rate = rc["limits"].get("max_yaw_rate", 60.0)
    ac = data.get("attitude_control") or {}
    for key, attr in (("roll", "roll_angle_gains"), ("pitch", "pitch_angle_gains")):
        if key in ac:
            setattr(config, attr, PIDGains.from_dict(ac[key]))
    if "limits" in ac:
        config.max_roll = ac["limits"].get("max_roll", 30.0)
        config.max_pitch = ac["limits"].get("max_pitch", 20.0)
    hc = data.get("hsa_control") or {}
    if "heading" in hc:
        config.hsa.heading_gains = PIDGains.from_dict(hc["heading"])
    if "limits" in hc:
        config.hsa.max_bank_angle = hc["limits"].get("max_bank_angle", 25.0)
        config.hsa.tecs.baseline_throttle = hc["limits"].get("baseline_throttle", 0.1)
        config.hsa.tecs.max_pitch_command = hc["limits"].get("max_pitch_command", 15.0)
    tecs = hc.get("tecs") or {}
    if "energy" in tecs:
        config.hsa.tecs.energy_gains = PIDGains.from_dict(tecs["energy"])
    if "balance" in tecs:
        config.hsa.tecs.balance_gains = PIDGains.from_dict(tecs["balance"])
    if "turn_compensation" in hc:
        config.hsa.tecs.load_factor_gain = hc["turn_compensation"].get("load_factor_gain", 0.05)
    wn = data.get("waypoint_navigation")
    if wn is not None:
        g = config.guidance
        g.acceptance_radius = wn.get("acceptance_radius", 40.0)
        pp, los, sc = wn.get("pure_pursuit"), wn.get("los"), wn.get("speed_control")
        if pp is not None:
            g.lookahead_time = pp.get("lookahead_time", 1.2)
            g.lookahead_min = pp.get("lookahead_min", 12.0)
            g.lookahead_max = pp.get("lookahead_max", 30.0)
            g.proximity_scale_distance = pp.get("proximity_scale_distance", 2.0)
        if los is not None:
            g.los_max_bank = los.get("max_bank_for_turn_calc", 20.0)
            g.los_lead_angle = los.get("lead_angle", 30.0)
        if sc is not None:
            g.turn_threshold_angle = sc.get("turn_threshold_angle", 60.0)
            g.turn_threshold_distance = sc.get("turn_threshold_distance", 60.0)
            g.max_speed_reduction = sc.get("max_speed_reduction", 0.15)
            g.min_speed = sc.get("min_speed", 14.0)
        if "proportional_nav" in wn:
            g.pn_gain = wn["proportional_nav"].get("gain", 3.0)
    return config


@dataclass
class MissionConfig:                               # config_loader.py:210-221
    name: str = "Unnamed Mission"
    pattern_type: str = "square"
    pattern_size: float = 300.0
    altitude: float = 100.0
    speed: float = 15.0
    guidance: str = "PP"
    max_duration: float = 180.0
    dt: float = 0.01
    output_dir: str = "final_figures"


def load_mission_config(config_file: str = "square_pattern.yaml", config_dir=None) -> MissionConfig:
    """config_loader.py:224-260"""
    filepath = _resolve(config_file, "missions", config_dir)
    config = MissionConfig()
    if not filepath.exists():
        print(f"Warning: Config file {filepath} not found, using defaults")
        return config
    data = load_yaml(filepath)
    m = data.get("mission")
    if m is not None:
        config.name = m.get("name", "Unnamed Mission")
        if "pattern" in m:
            config.pattern_type = m["pattern"].get("type", "square")
            config.pattern_size = m["pattern"].get("size", 300.0)
        if "flight" in m:
            config.altitude = m["flight"].get("altitude", 100.0)
            config.speed = m["flight"].get("speed", 15.0)
            config.guidance = m["flight"].get("guidance", "PP")
    s = data.get("simulation")
    if s is not None:
        config.max_duration = s.get("max_duration", 180.0)
        config.dt = s.get("dt", 0.01)
    if "output" in data:
        config.output_dir = data["output"].get("directory", "final_figures")
    return config


def square_mission_waypoints(mission: MissionConfig):
    """The square the waypoint demo flies for a MissionConfig (examples/03_waypoint_square_demo.py:120-135): corners of a
    ``pattern_size`` square at ``altitude``, every leg at ``speed``, back to the start.  Rows = north, east, down, speed."""
    s, d, v = float(mission.pattern_size), -float(mission.altitude), float(mission.speed)
    return [[0.0, 0.0, d, v], [s, 0.0, d, v], [s, s, d, v], [0.0, s, d, v], [0.0, 0.0, d, v]]


# ---- configuration -> acb200_controller_params ------------------------------------------------------------------
def _pid(gains, i_limit: float, omin: float, omax: float) -> _cabi.PidConfigC:
    """create_pid_config (controllers/utils/pid_utils.py:16-46): the pybind setters narrow every value to float."""
    return _cabi.PidConfigC(kp=gains.kp, ki=gains.ki, kd=gains.kd, output_min=omin, output_max=omax, integral_min=-i_limit,
                            integral_max=i_limit, derivative_filter_alpha=0.1)


def controller_params(config=None, flight_config: Optional[FlightControlConfig] = None) -> _cabi.ControllerParamsC:
    """What the reference's agent constructors make of ``(ControllerConfig, FlightControlConfig)``:

    * rate / angle loops and their limits come from the legacy ``ControllerConfig`` (rate_agent.py:41-55,
      attitude_agent.py:43-71) -- NOT from the YAML's rate_control / attitude_control sections, which the reference
      loads into ``FlightControlConfig`` but never hands to those agents (examples/03_waypoint_square_demo.py:109-117);
    * heading / TECS PIDs, bank and pitch limits, throttle baseline, load-factor gain from ``flight_config.hsa``
      (defaults ``HSAConfig()`` / ``TECSConfig()``), the heading integral limit from ``config.heading_gains.i_limit``
      (hsa_agent.py:47-108); the 10-degree pitch-command clamp is a constant there (:108);
    * guidance from ``flight_config.guidance`` (default ``GuidanceConfig()``), pure-pursuit bank from
      ``flight_config.hsa.max_bank_angle`` (default 25) (waypoint_agent.py:49-63)."""
    from .adapters import ControllerConfig
    c = config if config is not None else ControllerConfig()
    hsa = flight_config.hsa if flight_config is not None else HSAConfig()
    tecs = hsa.tecs
    g = flight_config.guidance if flight_config is not None else GuidanceConfig()
    rad = math.radians
    P = _cabi.ControllerParamsC()
    P.max_roll_rate, P.max_pitch_rate, P.max_yaw_rate = rad(c.max_roll_rate), rad(c.max_pitch_rate), rad(c.max_yaw_rate)
    P.max_roll, P.max_pitch = rad(c.max_roll), rad(c.max_pitch)
    for k, gains in enumerate((c.roll_rate_gains, c.pitch_rate_gains, c.yaw_gains)):
        P.rate[k] = _pid(gains, gains.i_limit, -1.0, 1.0)
    for k, (gains, lim) in enumerate(((c.roll_angle_gains, P.max_roll_rate), (c.pitch_angle_gains, P.max_pitch_rate),
                                      (c.yaw_gains, P.max_yaw_rate))):
        P.angle[k] = _pid(gains, gains.i_limit, -lim, lim)
    heading_i = getattr(getattr(c, "heading_gains", None), "i_limit", 25.0)
    P.max_bank = rad(hsa.max_bank_angle)
    P.heading = _pid(hsa.heading_gains, heading_i, -P.max_bank, P.max_bank)
    P.energy = _pid(tecs.energy_gains, 10.0, -0.5, 0.5)
    P.balance = _pid(tecs.balance_gains, 5.0, -rad(tecs.max_pitch_command), rad(tecs.max_pitch_command))
    P.baseline_throttle, P.load_factor_gain, P.max_pitch_cmd = tecs.baseline_throttle, tecs.load_factor_gain, rad(10.0)
    P.max_bank_angle_deg = hsa.max_bank_angle if flight_config is not None else 25.0
    P.lookahead_time, P.lookahead_min, P.lookahead_max = g.lookahead_time, g.lookahead_min, g.lookahead_max
    P.proximity_scale_distance = g.proximity_scale_distance
    P.los_max_bank_deg, P.los_lead_angle_deg = g.los_max_bank, g.los_lead_angle
    P.turn_threshold_angle_deg, P.turn_threshold_distance = g.turn_threshold_angle, g.turn_threshold_distance
    P.max_speed_reduction, P.min_speed = g.max_speed_reduction, g.min_speed
    return P
