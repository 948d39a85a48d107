"""ctypes front-end of oracle/libacb_oracle.so -- the CPU checker.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and the cpu_baseline / --impl reference legs of bench.py may
import this module.  The product package never does (tests/test_no_oracle_in_product.py enforces it).

All arrays are numpy, structure-of-arrays, field-major ``[field][n]`` (see acb_oracle.h).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libacb_oracle.so")

CTRL_EXTERNAL, CTRL_RATE_PID, CTRL_ATTITUDE_PID, CTRL_RESIDUAL = 0, 1, 2, 3
CMD_STEP, CMD_RAMP, CMD_SINE, CMD_RANDOM = 0, 1, 2, 3
OBS_DIM = 18

AIRCRAFT_FIELDS = [
    "mass", "inertia_xx", "inertia_yy", "inertia_zz", "wing_area", "wing_span", "chord",
    "cl_0", "cl_alpha", "cd_0", "cd_alpha2", "cl_elevator", "cm_elevator", "cy_rudder", "cn_rudder",
    "cl_aileron", "cm_alpha", "cn_beta", "cl_beta", "damping_roll", "damping_pitch", "damping_yaw",
    "max_thrust", "air_density", "gravity", "min_airspeed_aero", "min_u_velocity",
    "max_elevator_deflection", "max_aileron_deflection", "max_rudder_deflection", "thrust_zero_velocity",
    "max_velocity", "max_angular_rate", "max_pitch_angle", "max_alpha", "max_acceleration",
    "max_angular_acceleration", "max_timestep", "min_timestep",
]


class AircraftParams(C.Structure):
    _fields_ = [(f, C.c_double) for f in AIRCRAFT_FIELDS]


class PidConfig(C.Structure):
    _fields_ = [(f, C.c_float) for f in ("kp", "ki", "kd", "output_min", "output_max", "integral_min",
                                         "integral_max", "derivative_filter_alpha")]


class PidState(C.Structure):
    _fields_ = [(f, C.c_float) for f in ("error", "error_prev", "integral", "derivative",
                                         "derivative_filtered", "output")]


class ControllerParams(C.Structure):
    _fields_ = [
        ("rate", PidConfig * 3), ("angle", PidConfig * 3),
        ("max_roll_rate", C.c_double), ("max_pitch_rate", C.c_double), ("max_yaw_rate", C.c_double),
        ("max_roll", C.c_double), ("max_pitch", C.c_double),
        ("heading", PidConfig), ("energy", PidConfig), ("balance", PidConfig),
        ("max_bank_rad", C.c_double), ("baseline_throttle", C.c_double), ("load_factor_gain", C.c_double),
        ("max_pitch_cmd_rad", C.c_double),
        ("max_bank_angle_deg", C.c_double), ("lookahead_time", C.c_double), ("lookahead_min", C.c_double),
        ("lookahead_max", C.c_double), ("proximity_scale_distance", C.c_double), ("los_max_bank_deg", C.c_double),
        ("los_lead_angle_deg", C.c_double), ("turn_threshold_angle_deg", C.c_double),
        ("turn_threshold_distance", C.c_double), ("max_speed_reduction", C.c_double), ("min_speed", C.c_double),
    ]


class EnvBuffers(C.Structure):
    _fields_ = [
        ("n", C.c_int64),
        ("state", C.c_void_p), ("derived", C.c_void_p), ("pid", C.c_void_p), ("att_cmd", C.c_void_p),
        ("rate_cmd", C.c_void_p), ("prev_action", C.c_void_p), ("rew_mem", C.c_void_p), ("sched", C.c_void_p),
        ("env_time", C.c_void_p), ("step_count", C.c_void_p), ("is_settled", C.c_void_p),
        ("command_type", C.c_void_p), ("difficulty", C.c_void_p), ("episode_idx", C.c_void_p),
        ("ep_return", C.c_void_p),
        ("obs", C.c_void_p), ("reward", C.c_void_p), ("terminated", C.c_void_p), ("truncated", C.c_void_p),
        ("terminal_obs", C.c_void_p), ("components", C.c_void_p), ("stats", C.c_void_p),
        ("walk_z", C.c_void_p),      # oracle only: injected standard normals [3][n] of the random-walk command
    ]


class EnvConfig(C.Structure):
    _fields_ = [
        ("dt", C.c_double), ("dt_physics", C.c_double), ("max_steps", C.c_int32), ("controller", C.c_int32),
        ("auto_reset", C.c_int32), ("pid_throttle", C.c_double), ("residual_scale", C.c_double),
        ("seed", C.c_uint64), ("id_offset", C.c_int64),
    ]


SENSOR_STDDEV_KEYS = ("imu_gyro_stddev", "imu_accel_stddev", "gps_position_stddev", "gps_velocity_stddev", "airspeed_stddev",
                      "altitude_stddev", "attitude_stddev")
SENSOR_DEFAULTS = dict(zip(SENSOR_STDDEV_KEYS, (0.01, 0.1, 1.0, 0.1, 0.5, 0.5, 0.01)))     # interfaces/sensor.py:176-182


class SensorConfig(C.Structure):
    _fields_ = [("enabled", C.c_int32), ("_pad", C.c_int32)] + [(k, C.c_double) for k in SENSOR_STDDEV_KEYS] + [
        ("gyro_bias_walk_stddev", C.c_double), ("accel_bias_walk_stddev", C.c_double), ("seed", C.c_uint64),
        ("id_offset", C.c_int64)]


def build(force: bool = False) -> str:
    """Compile the restatement (and nothing else) if the .so is missing or stale."""
    src = os.path.join(_HERE, "acb_oracle.c")
    hdr = os.path.join(_HERE, "acb_oracle.h")
    stale = (not os.path.exists(_LIB_PATH)
             or os.path.getmtime(_LIB_PATH) < max(os.path.getmtime(src), os.path.getmtime(hdr)))
    if force or stale:
        subprocess.run(["make", "-C", _HERE, "libacb_oracle.so"], check=True, capture_output=True)
    return _LIB_PATH


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        L.orc_default_aircraft_params.argtypes = [C.POINTER(AircraftParams)]
        L.orc_default_controller_params.argtypes = [C.POINTER(ControllerParams)]
        L.orc_dynamics.argtypes = [C.POINTER(AircraftParams), C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_rk4_step.argtypes = [C.POINTER(AircraftParams), C.c_void_p, C.c_void_p, C.c_void_p, C.c_double]
        L.orc_derived.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_pid_compute.argtypes = [C.POINTER(PidConfig), C.POINTER(PidState), C.c_float, C.c_float, C.c_float]
        L.orc_pid_compute.restype = C.c_float
        L.orc_wrap_angle.argtypes = [C.c_double]
        L.orc_wrap_angle.restype = C.c_double
        L.orc_step_surface.argtypes = [C.POINTER(AircraftParams), C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.c_double, C.c_double, C.c_int]
        L.orc_step_cascade.argtypes = [C.POINTER(AircraftParams), C.POINTER(ControllerParams), C.c_int, C.c_int64,
                                       C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.c_double, C.c_double, C.c_double, C.c_int]
        L.orc_step_hsa.argtypes = [C.POINTER(AircraftParams), C.POINTER(ControllerParams), C.c_int64, C.c_void_p, C.c_void_p,
                                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_int]
        L.orc_step_waypoint.argtypes = [C.POINTER(AircraftParams), C.POINTER(ControllerParams), C.c_int64, C.c_void_p,
                                        C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_double, C.c_int, C.c_void_p, C.c_void_p,
                                        C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_int]
        L.orc_step_waypoint_stats.argtypes = L.orc_step_waypoint.argtypes + [C.c_void_p]
        L.orc_env_reset.argtypes = [C.POINTER(AircraftParams), C.POINTER(EnvConfig), C.POINTER(EnvBuffers),
                                    C.c_void_p, C.c_int]
        L.orc_env_step.argtypes = [C.POINTER(AircraftParams), C.POINTER(ControllerParams), C.POINTER(EnvConfig),
                                   C.POINTER(EnvBuffers), C.c_void_p, C.c_int]
        L.orc_philox4x32_10.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_walk_normals.argtypes = [C.c_uint64, C.c_int64, C.c_uint32, C.c_int32, C.c_void_p]
        L.orc_difficulty_scale.argtypes = [C.c_float]
        L.orc_difficulty_scale.restype = C.c_double
        L.orc_sensor_normals.argtypes = [C.c_uint64, C.c_int64, C.c_uint32, C.c_void_p]
        L.orc_sensor_update.argtypes = [C.POINTER(SensorConfig), C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32,
                                        C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def default_aircraft_params() -> AircraftParams:
    p = AircraftParams()
    lib().orc_default_aircraft_params(C.byref(p))
    return p


def default_controller_params() -> ControllerParams:
    c = ControllerParams()
    lib().orc_default_controller_params(C.byref(c))
    return c


def _ptr(a):
    return None if a is None else a.ctypes.data


def _chk(a, dtype, shape=None):
    assert a.dtype == dtype and a.flags.c_contiguous, (a.dtype, dtype)
    if shape is not None:
        assert a.shape == tuple(shape), (a.shape, shape)
    return a


def dynamics(params, s, ctrl):
    out = np.zeros(12)
    lib().orc_dynamics(C.byref(params), _chk(s, np.float64, (12,)).ctypes.data,
                       _chk(ctrl, np.float64, (4,)).ctypes.data, out.ctypes.data)
    return out


def derived(s):
    out = np.zeros(4)
    lib().orc_derived(_chk(s, np.float64, (12,)).ctypes.data, out.ctypes.data)
    return out


def wrap_angle(a: float) -> float:
    return lib().orc_wrap_angle(a)


class Pid:
    """One scalar PID (cpp PIDController) on the restated core."""

    def __init__(self, cfg: PidConfig):
        self.cfg = cfg
        self.st = PidState()

    def compute(self, sp, meas, dt) -> float:
        return lib().orc_pid_compute(C.byref(self.cfg), C.byref(self.st), sp, meas, dt)


def step_surface(params, state, ctrl, dt, dt_physics=0.001, ctrl_held=None, derived_out=None, nthreads=0) -> int:
    n = state.shape[1]
    _chk(state, np.float64, (13, n)); _chk(ctrl, np.float64, (4, n))
    if ctrl_held is not None: _chk(ctrl_held, np.float64, (4, n))
    if derived_out is not None: _chk(derived_out, np.float64, (4, n))
    k = lib().orc_step_surface(C.byref(params), n, state.ctypes.data, ctrl.ctypes.data, _ptr(ctrl_held),
                               _ptr(derived_out), dt, dt_physics, nthreads)
    if k < 0:
        raise ValueError(f"Invalid timestep dt={dt}")
    return k


def step_cascade(params, cparams, level, state, cmd, pid, dt, dt_ctrl=None, dt_physics=0.001,
                 surfaces_out=None, derived_out=None, nthreads=0) -> int:
    n = state.shape[1]
    _chk(state, np.float64, (13, n)); _chk(cmd, np.float64, (4, n)); _chk(pid, np.float32, (18, n))
    if surfaces_out is not None: _chk(surfaces_out, np.float64, (4, n))
    if derived_out is not None: _chk(derived_out, np.float64, (4, n))
    k = lib().orc_step_cascade(C.byref(params), C.byref(cparams), level, n, state.ctypes.data, cmd.ctypes.data,
                               pid.ctypes.data, _ptr(surfaces_out), _ptr(derived_out),
                               dt if dt_ctrl is None else dt_ctrl, dt, dt_physics, nthreads)
    if k < 0:
        raise ValueError(f"step_cascade failed ({k})")
    return k


GUIDANCE_LOS, GUIDANCE_PP = 0, 1


def step_hsa(params, cparams, state, cmd, pid, pid_outer, dt, dt_ctrl=None, dt_physics=0.001, surfaces_out=None,
             derived_out=None, nthreads=0) -> int:
    """HSAAgent.compute_action + set_controls + step; cmd [3][n] = heading, speed, altitude."""
    n = state.shape[1]
    _chk(state, np.float64, (13, n)); _chk(cmd, np.float64, (3, n)); _chk(pid, np.float32, (18, n)); _chk(pid_outer, np.float32, (9, n))
    k = lib().orc_step_hsa(C.byref(params), C.byref(cparams), n, state.ctypes.data, cmd.ctypes.data, pid.ctypes.data,
                           pid_outer.ctypes.data, _ptr(surfaces_out), _ptr(derived_out), dt if dt_ctrl is None else dt_ctrl, dt,
                           dt_physics, nthreads)
    if k < 0:
        raise ValueError(f"step_hsa failed ({k})")
    return k


def step_waypoint(params, cparams, state, waypoints, wp_index, pid, pid_outer, dt, guidance=GUIDANCE_PP, acceptance_radius=40.0,
                  loop=False, dt_ctrl=None, dt_physics=0.001, surfaces_out=None, derived_out=None, hsa_cmd_out=None,
                  nthreads=0, mission_stats=None) -> int:
    """MissionPlanner.update + WaypointAgent.compute_action + step; waypoints [n_wp][4] = north, east, down, speed.
    mission_stats: float64[16] accumulated (the fields of acb200_mission.stats)."""
    n = state.shape[1]
    _chk(state, np.float64, (13, n)); _chk(waypoints, np.float64); _chk(wp_index, np.int32, (n,))
    _chk(pid, np.float32, (18, n)); _chk(pid_outer, np.float32, (9, n))
    assert waypoints.ndim == 2 and waypoints.shape[1] == 4
    if mission_stats is not None:
        _chk(mission_stats, np.float64, (16,))
    k = lib().orc_step_waypoint_stats(C.byref(params), C.byref(cparams), n, state.ctypes.data, waypoints.ctypes.data,
                                      waypoints.shape[0], wp_index.ctypes.data, guidance, acceptance_radius, int(loop),
                                      pid.ctypes.data, pid_outer.ctypes.data, _ptr(surfaces_out), _ptr(derived_out),
                                      _ptr(hsa_cmd_out), dt if dt_ctrl is None else dt_ctrl, dt, dt_physics, nthreads,
                                      _ptr(mission_stats))
    if k < 0:
        raise ValueError(f"step_waypoint failed ({k})")
    return k


class Env:
    """The restated RateControlEnv over n aircraft (numpy SoA buffers owned here)."""

    def __init__(self, n, controller=CTRL_EXTERNAL, dt=0.02, episode_length=10.0, auto_reset=True, seed=0,
                 id_offset=0, difficulty=0.5, command_type=CMD_STEP, pid_throttle=0.6, residual_scale=0.3,
                 params=None, cparams=None, nthreads=0):
        self.n = n
        self.params = params or default_aircraft_params()
        self.cparams = cparams or default_controller_params()
        self.nthreads = nthreads
        self.cfg = EnvConfig(dt=dt, dt_physics=0.001, max_steps=int(episode_length / dt), controller=controller,
                             auto_reset=int(auto_reset), pid_throttle=pid_throttle, residual_scale=residual_scale,
                             seed=seed, id_offset=id_offset)
        f8, f4 = np.float64, np.float32
        self.state = np.zeros((13, n), f8); self.derived = np.zeros((4, n), f8)
        self.pid = np.zeros((18, n), f4); self.att_cmd = np.zeros((4, n), f8)
        self.rate_cmd = np.zeros((3, n), f8); self.prev_action = np.zeros((4, n), f4)
        self.rew_mem = np.zeros((7, n), f8); self.sched = np.zeros((7, n), f8)
        self.env_time = np.zeros(n, f8); self.step_count = np.zeros(n, np.int32)
        self.is_settled = np.zeros(n, np.uint8)
        self.command_type = np.full(n, command_type, np.uint8) if np.isscalar(command_type) else np.asarray(command_type, np.uint8).copy()
        self.difficulty = np.full(n, difficulty, f4) if np.isscalar(difficulty) else np.asarray(difficulty, f4).copy()
        self.episode_idx = np.zeros(n, np.uint32); self.ep_return = np.zeros(n, f8)
        self.obs = np.zeros((n, OBS_DIM), f4); self.reward = np.zeros(n, f8)
        self.terminated = np.zeros(n, np.uint8); self.truncated = np.zeros(n, np.uint8)
        self.terminal_obs = np.zeros((n, OBS_DIM), f4); self.components = np.zeros((8, n), f8)
        self.stats = np.zeros(16, f8)
        self.buf = EnvBuffers(n=n)
        for name, _ in EnvBuffers._fields_[1:]:
            if name != "walk_z":
                setattr(self.buf, name, getattr(self, name).ctypes.data)
        self._walk_z = None

    def reset(self, mask=None):
        m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
        lib().orc_env_reset(C.byref(self.params), C.byref(self.cfg), C.byref(self.buf), _ptr(m), self.nthreads)
        return self.obs

    def step(self, action=None, walk_z=None):
        """walk_z [3][n] float64: standard normals that replace this step's Philox draw of the random-walk command
        (the reference's own ``randn(3)`` values, replayed)."""
        a = None if action is None else _chk(np.ascontiguousarray(action, np.float32), np.float32, (self.n, 4))
        self._walk_z = None if walk_z is None else _chk(np.ascontiguousarray(walk_z, np.float64), np.float64, (3, self.n))
        self.buf.walk_z = _ptr(self._walk_z)
        k = lib().orc_env_step(C.byref(self.params), C.byref(self.cparams), C.byref(self.cfg), C.byref(self.buf),
                               _ptr(a), self.nthreads)
        if k < 0:
            raise ValueError("Invalid timestep")
        return self.obs, self.reward, self.terminated, self.truncated


def walk_normals(seed: int, global_id: int, episode: int, step_count: int) -> np.ndarray:
    z = np.zeros(3)
    lib().orc_walk_normals(seed, global_id, episode, step_count, z.ctypes.data)
    return z


def difficulty_scale(d: float) -> float:
    return lib().orc_difficulty_scale(d)


def philox(ctr, key):
    c = np.asarray(ctr, np.uint32); k = np.asarray(key, np.uint32); out = np.zeros(4, np.uint32)
    lib().orc_philox4x32_10(c.ctypes.data, k.ctypes.data, out.ctypes.data)
    return out


# ---- NoisySensorInterface (interfaces/sensor.py:137-263) ---------------------------------------------------------------
def sensor_config(noise_config: dict | None = None, id_offset: int = 0) -> SensorConfig:
    """The reference's noise_config dict (sensor.py:146-182) -> POD.  ``seed`` None is not reproducible there; 0 here."""
    nc = dict(noise_config or {})
    c = SensorConfig(enabled=int(bool(nc.get("enabled", True))), gyro_bias_walk_stddev=0.0001, accel_bias_walk_stddev=0.001,
                     seed=int(nc.get("seed") or 0), id_offset=id_offset)
    for k in SENSOR_STDDEV_KEYS:
        setattr(c, k, float(nc.get(k, SENSOR_DEFAULTS[k])))
    return c


def sensor_normals(seed: int, global_id: int, tick: int) -> np.ndarray:
    z = np.zeros(20)
    lib().orc_sensor_normals(seed, global_id, tick, z.ctypes.data)
    return z


def sensor_update(cfg: SensorConfig, state, derived_in, bias, tick: int, z=None):
    """state [13][n], derived [4][n] float64 (truth); bias [6][n] float64 advanced in place -> (m_state, m_derived).
    z [20][n]: caller-supplied standard normals in the reference's draw order instead of the Philox stream."""
    n = state.shape[1]
    _chk(state, np.float64, (13, n)); _chk(derived_in, np.float64, (4, n)); _chk(bias, np.float64, (6, n))
    if z is not None:
        _chk(z, np.float64, (20, n))
    ms, md = np.empty_like(state), np.empty_like(derived_in)
    if lib().orc_sensor_update(C.byref(cfg), n, _ptr(state), _ptr(derived_in), _ptr(bias), tick, _ptr(z), _ptr(ms),
                               _ptr(md)) != 0:
        raise ValueError("scale < 0")
    return ms, md
