"""ctypes binding of libacb200.so (include/acb200.h).  No CPU fallback: if the library or a CUDA
device is missing, everything here raises."""
from __future__ import annotations

import ctypes as C
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ACB200_LIB") or os.path.join(_PKG, "libacb200.so")   # ACB200_LIB: A/B builds of the same ABI

OK, EVALUE, ECUDA, ENULL, ENOSUP = 0, -1, -2, -3, -4
LEVEL_WAYPOINT, LEVEL_HSA, LEVEL_ATTITUDE, LEVEL_RATE, LEVEL_SURFACE = 1, 2, 3, 4, 5
CTRL_EXTERNAL, CTRL_RATE_PID, CTRL_ATTITUDE_PID, CTRL_RESIDUAL = 0, 1, 2, 3
CMD_STEP, CMD_RAMP, CMD_SINE, CMD_RANDOM = 0, 1, 2, 3
OBS_DIM, STATE_FIELDS, CTRL_FIELDS, DERIVED_FIELDS, PID_FIELDS, STATS_FIELDS = 18, 13, 4, 4, 18, 16
OUTER_PID_FIELDS = 9
GUIDANCE_LOS, GUIDANCE_PP = 0, 1

AIRCRAFT_FIELDS = [
    "mass", "inertia_xx", "inertia_yy", "inertia_zz", "wing_area", "wing_span", "chord",
    "cl_0", "cl_alpha", "cd_0", "cd_alpha2", "cl_elevator", "cm_elevator", "cy_rudder", "cn_rudder",
    "cl_aileron", "cm_alpha", "cn_beta", "cl_beta", "damping_roll", "damping_pitch", "damping_yaw",
    "max_thrust", "air_density", "gravity", "min_airspeed_aero", "min_u_velocity",
    "max_elevator_deflection", "max_aileron_deflection", "max_rudder_deflection", "thrust_zero_velocity",
    "max_velocity", "max_angular_rate", "max_pitch_angle", "max_alpha", "max_acceleration",
    "max_angular_acceleration", "max_timestep", "min_timestep",
]


class AircraftParamsC(C.Structure):
    _fields_ = [(f, C.c_double) for f in AIRCRAFT_FIELDS]


class PidConfigC(C.Structure):
    _fields_ = [(f, C.c_float) for f in ("kp", "ki", "kd", "output_min", "output_max", "integral_min",
                                         "integral_max", "derivative_filter_alpha")]


class PidStateC(C.Structure):
    _fields_ = [(f, C.c_float) for f in ("error", "error_prev", "integral", "derivative",
                                         "derivative_filtered", "output")]


class ControllerParamsC(C.Structure):
    _fields_ = [("rate", PidConfigC * 3), ("angle", PidConfigC * 3),
                ("max_roll_rate", C.c_double), ("max_pitch_rate", C.c_double), ("max_yaw_rate", C.c_double),
                ("max_roll", C.c_double), ("max_pitch", C.c_double),
                ("heading", PidConfigC), ("energy", PidConfigC), ("balance", PidConfigC),
                ("max_bank", C.c_double), ("baseline_throttle", C.c_double), ("load_factor_gain", C.c_double),
                ("max_pitch_cmd", C.c_double), ("max_bank_angle_deg", C.c_double), ("lookahead_time", C.c_double),
                ("lookahead_min", C.c_double), ("lookahead_max", C.c_double), ("proximity_scale_distance", C.c_double),
                ("los_max_bank_deg", C.c_double), ("los_lead_angle_deg", C.c_double),
                ("turn_threshold_angle_deg", C.c_double), ("turn_threshold_distance", C.c_double),
                ("max_speed_reduction", C.c_double), ("min_speed", C.c_double)]


SENSOR_STDDEV_KEYS = ("imu_gyro_stddev", "imu_accel_stddev", "gps_position_stddev", "gps_velocity_stddev", "airspeed_stddev",
                      "altitude_stddev", "attitude_stddev")


class SensorConfigC(C.Structure):
    _fields_ = [("enabled", C.c_int32), ("_pad", C.c_int32)] + [(k, C.c_double) for k in SENSOR_STDDEV_KEYS] + [
        ("gyro_bias_walk_stddev", C.c_double), ("accel_bias_walk_stddev", C.c_double), ("seed", C.c_uint64),
        ("id_offset", C.c_int64)]


class MissionC(C.Structure):
    _fields_ = [("waypoints", C.c_void_p), ("n_waypoints", C.c_int32), ("guidance", C.c_int32), ("wp_index", C.c_void_p),
                ("loop", C.c_int32), ("_pad", C.c_int32), ("acceptance_radius", C.c_double), ("stats", C.c_void_p)]


MISSION_STATS_NAMES = ("aircraft_steps", "waypoints_reached", "laps", "missions_completed", "cross_track_abs_sum", "cross_track_sq_sum",
                       "altitude_error_abs_sum", "distance_to_waypoint_sum", "airspeed_sum", "steps_below_5m")


class FleetC(C.Structure):
    _fields_ = [("n", C.c_int64), ("precision", C.c_int32), ("_pad", C.c_int32),
                ("state", C.c_void_p), ("controls", C.c_void_p), ("derived", C.c_void_p), ("pid", C.c_void_p)]


ENV_POINTER_FIELDS = ["att_cmd", "rate_cmd", "prev_action", "rew_mem", "sched", "env_time", "step_count",
                      "is_settled", "command_type", "difficulty", "episode_idx", "ep_return", "obs", "reward",
                      "terminated", "truncated", "terminal_obs", "components", "stats"]


class EnvC(C.Structure):
    _fields_ = [("fleet", FleetC)] + [(f, C.c_void_p) for f in ENV_POINTER_FIELDS]


class EnvConfigC(C.Structure):
    _fields_ = [("dt", C.c_double), ("dt_physics", C.c_double), ("max_steps", C.c_int32), ("controller", C.c_int32),
                ("auto_reset", C.c_int32), ("_pad", C.c_int32), ("pid_throttle", C.c_double),
                ("residual_scale", C.c_double), ("seed", C.c_uint64), ("id_offset", C.c_int64)]


_FP = C.POINTER(C.c_float)


class LstmPolicyWeightsC(C.Structure):
    _fields_ = [(k, _FP) for k in (
        "embed_w", "embed_b", "l0_w_ih", "l0_b_ih", "l0_b_hh", "l1_w_ih", "l1_b_ih", "l1_b_hh", "proj_w", "proj_b",
        "actor_w_ih", "actor_w_hh", "actor_b_ih", "actor_b_hh", "pi0_w", "pi0_b", "pi1_w", "pi1_b", "action_w", "action_b")]


class MlpPolicyWeightsC(C.Structure):
    _fields_ = [(k, _FP) for k in ("pi0_w", "pi0_b", "pi1_w", "pi1_b", "pi2_w", "pi2_b", "action_w", "action_b")]


# every symbol include/acb200.h declares (tests/test_cabi_symbols.py checks the header against this)
SYMBOLS = [
    "acb200_version", "acb200_create", "acb200_destroy", "acb200_last_error", "acb200_set_stream",
    "acb200_synchronize", "acb200_default_aircraft_params", "acb200_cessna_aircraft_params",
    "acb200_default_controller_params", "acb200_set_aircraft_params", "acb200_set_controller_params",
    "acb200_fleet_reset", "acb200_step_surface", "acb200_step_cascade", "acb200_step_hsa", "acb200_step_waypoint", "acb200_env_reset", "acb200_env_step",
    "acb200_env_step_host", "acb200_set_host_chunks", "acb200_pid_compute_f32", "acb200_pid_reset", "acb200_kernel_info",
    "acb200_launch_count", "acb200_measure_fp32_peak",
    "acb200_policy_create_lstm", "acb200_policy_create_mlp", "acb200_policy_destroy", "acb200_policy_state_bytes",
    "acb200_policy_forward", "acb200_policy_state_export", "acb200_policy_state_import", "acb200_policy_info",
    "acb200_gae",
    "acb200_env_step_range", "acb200_telemetry_record",
    "acb200_default_sensor_config", "acb200_sensor_update", "acb200_sensor_reset",
    "acb200_validate_aircraft_params", "acb200_get_aircraft_params", "acb200_get_controller_params",
    "acb200_set_param_groups", "acb200_env_gather_done",
    "acb200_env_gather_done_noalloc", "acb200_noalloc_max_aircraft",
    "acb200_param_table_words", "acb200_build_param_table", "acb200_set_param_table",
    "acb200_policy_env_rollout",
    "acb200_comm_unique_id", "acb200_comm_init", "acb200_comm_destroy", "acb200_comm_info", "acb200_episode_stats",
]

_lib = None


def load() -> C.CDLL:
    """dlopen the in-tree library.  Raises if it has not been built (python __graft_entry__.py build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, dbl, flt = C.c_void_p, C.c_int, C.c_int64, C.c_double, C.c_float
    P = C.POINTER
    L.acb200_version.restype = C.c_char_p
    L.acb200_create.argtypes = [i32, P(vp)]
    L.acb200_destroy.argtypes = [vp]; L.acb200_destroy.restype = None
    L.acb200_last_error.argtypes = [vp]; L.acb200_last_error.restype = C.c_char_p
    L.acb200_set_stream.argtypes = [vp, vp]
    L.acb200_synchronize.argtypes = [vp]
    L.acb200_default_aircraft_params.argtypes = [P(AircraftParamsC)]; L.acb200_default_aircraft_params.restype = None
    L.acb200_cessna_aircraft_params.argtypes = [P(AircraftParamsC)]; L.acb200_cessna_aircraft_params.restype = None
    L.acb200_default_controller_params.argtypes = [P(ControllerParamsC)]; L.acb200_default_controller_params.restype = None
    L.acb200_set_aircraft_params.argtypes = [vp, P(AircraftParamsC)]
    L.acb200_set_controller_params.argtypes = [vp, P(ControllerParamsC)]
    L.acb200_validate_aircraft_params.argtypes = [vp, P(AircraftParamsC)]
    L.acb200_get_aircraft_params.argtypes = [vp, P(AircraftParamsC)]
    L.acb200_get_controller_params.argtypes = [vp, P(ControllerParamsC)]
    L.acb200_set_param_groups.argtypes = [vp, vp, P(i64), C.c_int32]
    L.acb200_env_gather_done.argtypes = [vp, P(EnvC), i64, vp, vp, vp, vp]
    L.acb200_policy_env_rollout.argtypes = [vp, vp, P(EnvC), P(EnvConfigC), vp, vp, vp, C.c_int32, C.c_int32]
    L.acb200_comm_unique_id.argtypes = [vp, vp]
    L.acb200_comm_init.argtypes = [vp, i32, i32, vp]
    L.acb200_comm_destroy.argtypes = [vp]
    L.acb200_comm_info.argtypes = [vp, P(i32), P(i32), P(i32)]
    L.acb200_episode_stats.argtypes = [vp, vp, vp, C.c_int32, C.c_int32]
    L.acb200_param_table_words.argtypes = []
    L.acb200_build_param_table.argtypes = [vp, vp, i64, i64, i64, C.c_int32, vp]
    L.acb200_set_param_table.argtypes = [vp, vp, i64, C.c_int32]
    L.acb200_env_gather_done_noalloc.argtypes = [vp, P(EnvC), i64, vp, vp, vp, vp]
    L.acb200_noalloc_max_aircraft.argtypes = []; L.acb200_noalloc_max_aircraft.restype = i64
    L.acb200_fleet_reset.argtypes = [vp, P(FleetC), vp, vp]
    L.acb200_step_surface.argtypes = [vp, P(FleetC), vp, dbl, dbl]
    L.acb200_step_cascade.argtypes = [vp, P(FleetC), i32, vp, dbl, dbl, dbl, vp]
    L.acb200_step_hsa.argtypes = [vp, P(FleetC), vp, vp, dbl, dbl, dbl, vp]
    L.acb200_step_waypoint.argtypes = [vp, P(FleetC), vp, P(MissionC), dbl, dbl, dbl, vp, vp]
    L.acb200_env_reset.argtypes = [vp, P(EnvC), P(EnvConfigC), vp]
    L.acb200_env_step.argtypes = [vp, P(EnvC), P(EnvConfigC), vp]
    L.acb200_env_step_range.argtypes = [vp, P(EnvC), P(EnvConfigC), vp, i64, i64]
    L.acb200_env_step_host.argtypes = [vp, P(EnvC), P(EnvConfigC), vp, vp, vp, vp, vp, vp, vp]
    L.acb200_set_host_chunks.argtypes = [vp, i32]
    L.acb200_pid_compute_f32.argtypes = [P(PidConfigC), P(PidStateC), flt, flt, flt]; L.acb200_pid_compute_f32.restype = flt
    L.acb200_pid_reset.argtypes = [P(PidStateC)]; L.acb200_pid_reset.restype = None
    L.acb200_kernel_info.argtypes = [i32, C.c_char_p, P(i32), P(i32), P(i32)]
    L.acb200_launch_count.argtypes = [vp]; L.acb200_launch_count.restype = i64
    L.acb200_measure_fp32_peak.argtypes = [vp, P(dbl)]
    L.acb200_policy_create_lstm.argtypes = [vp, P(LstmPolicyWeightsC), P(vp)]
    L.acb200_policy_create_mlp.argtypes = [vp, P(MlpPolicyWeightsC), P(vp)]
    L.acb200_policy_destroy.argtypes = [vp]; L.acb200_policy_destroy.restype = None
    L.acb200_policy_state_bytes.argtypes = [vp, i64, P(C.c_size_t), P(C.c_size_t)]
    L.acb200_policy_forward.argtypes = [vp, vp, i64, vp, vp, vp, vp, vp, vp]
    L.acb200_policy_state_export.argtypes = [vp, vp, i64, vp, vp, vp, vp]
    L.acb200_policy_state_import.argtypes = [vp, vp, i64, vp, vp, vp, vp]
    L.acb200_policy_info.argtypes = [vp, P(dbl), P(C.c_size_t), P(i32)]
    L.acb200_gae.argtypes = [vp, C.c_int32, i64, vp, vp, vp, vp, vp, dbl, dbl, vp, vp]
    L.acb200_default_sensor_config.argtypes = [P(SensorConfigC)]
    L.acb200_default_sensor_config.restype = None
    L.acb200_sensor_update.argtypes = [vp, P(FleetC), P(SensorConfigC), vp, C.c_uint32, vp, vp]
    L.acb200_sensor_reset.argtypes = [vp, i64, i32, vp, vp]
    L.acb200_telemetry_record.argtypes = [vp, P(FleetC), vp, i64, vp]
    _lib = L
    return L


class Context:
    """acb200_ctx: one per device.  Maps error codes to the reference's exception types."""

    def __init__(self, device: int = 0):
        self.lib = load()
        h = C.c_void_p()
        rc = self.lib.acb200_create(device, C.byref(h))
        if rc != OK:
            raise RuntimeError(f"acb200_create(device={device}) failed ({rc}): no usable CUDA device "
                               "(this library has no CPU path)")
        self.handle = h
        self.device = device

    def close(self):
        if getattr(self, "handle", None):
            self.lib.acb200_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def check(self, rc: int):
        if rc == OK:
            return
        msg = self.lib.acb200_last_error(self.handle).decode()
        if rc == EVALUE:
            raise ValueError(msg)
        raise RuntimeError(f"acb200 error {rc}: {msg}")

    def set_stream(self, cuda_stream: int):
        self.check(self.lib.acb200_set_stream(self.handle, C.c_void_p(cuda_stream)))

    def synchronize(self):
        self.check(self.lib.acb200_synchronize(self.handle))

    def measure_fp32_peak(self) -> float:
        """TFLOP/s of an FMA-chain kernel on this device (non-tensor FP32 roofline denominator)."""
        v = C.c_double()
        self.check(self.lib.acb200_measure_fp32_peak(self.handle, C.byref(v)))
        return v.value

    @property
    def launch_count(self) -> int:
        return self.lib.acb200_launch_count(self.handle)


def default_aircraft_params(kind: str = "rc_plane") -> AircraftParamsC:
    p = AircraftParamsC()
    (load().acb200_cessna_aircraft_params if kind == "cessna" else load().acb200_default_aircraft_params)(C.byref(p))
    return p


def default_controller_params() -> ControllerParamsC:
    c = ControllerParamsC()
    load().acb200_default_controller_params(C.byref(c))
    return c


def kernel_info(precision: int, name: str):
    r, l, m = C.c_int(), C.c_int(), C.c_int()
    rc = load().acb200_kernel_info(precision, name.encode(), C.byref(r), C.byref(l), C.byref(m))
    if rc != OK:
        raise RuntimeError(f"kernel_info({name}) failed: {rc}")
    return {"regs": r.value, "local_bytes": l.value, "max_threads": m.value}
