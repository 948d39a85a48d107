"""Batched aircraft backend: N Simplified6DOF simulators as one structure-of-arrays on the GPU.

Mirrors SimulationAircraftBackend (simulation/simulation_backend.py:13-170 in the reference) with
tensors instead of dataclasses: ``set_controls`` / ``step(dt)`` / ``reset`` / ``get_state`` keep their
meaning; the cascaded agents (``step_rate``, ``step_attitude``) are fused with the physics step.
torch is only the allocator and stream provider here; all arithmetic happens in libacb200.so.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _cabi


def _dtype(precision: int):
    if precision == 64:
        return torch.float64
    if precision == 32:
        return torch.float32
    raise ValueError("precision must be 64 (parity build) or 32 (throughput build)")


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def as_tensor(x, device, dtype=None) -> torch.Tensor:
    """Tensor intake of the adapters: a torch tensor, or ANY object with ``__dlpack__`` (CuPy, JAX, numba device arrays,
    another framework's CUDA buffer -- consumed zero-copy through ``torch.from_dlpack``, north_star's "exchanges DLPack /
    PyTorch tensors"), or a numpy array / nested list (copied).  The library itself only ever sees ``data_ptr()``."""
    if isinstance(x, torch.Tensor):
        t = x
    elif hasattr(x, "__dlpack__"):
        t = torch.from_dlpack(x)
    else:
        t = torch.as_tensor(x)
    if t.device != device or (dtype is not None and t.dtype != dtype):
        t = t.to(device=device, dtype=dtype if dtype is not None else t.dtype)
    return t.contiguous()


class AircraftFleet:
    """N aircraft.  state[13][n] = N,E,D,u,v,w,roll,pitch,yaw,p,q,r,time; controls[4][n] = elevator,
    aileron, rudder, throttle; derived[4][n] = airspeed, altitude, ground_speed, heading; pid[18][n]."""

    def __init__(self, n: int, precision: int = 64, device: int | str | torch.device = 0, config: Optional[dict] = None,
                 ctx: Optional[_cabi.Context] = None):
        if not torch.cuda.is_available():
            raise RuntimeError("acb200 needs a CUDA device (no CPU fallback)")
        self.device = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
        self.n, self.precision, self.dtype = int(n), precision, _dtype(precision)
        config = config or {}
        self.dt_physics = float(config.get("dt_physics", 0.001))          # simulation_backend.py:50
        self.ctx = ctx or _cabi.Context(self.device.index or 0)
        self.aircraft_params = config.get("params") or _cabi.default_aircraft_params(config.get("aircraft_type", "rc_plane"))
        self.ctx.check(self.ctx.lib.acb200_set_aircraft_params(self.ctx.handle, C.byref(self.aircraft_params)))
        self.controller_params = config.get("controller_params") or _cabi.default_controller_params()
        self.ctx.check(self.ctx.lib.acb200_set_controller_params(self.ctx.handle, C.byref(self.controller_params)))
        kw = dict(device=self.device, dtype=self.dtype)
        self.state = torch.zeros((13, n), **kw)
        self.controls = torch.zeros((4, n), **kw)
        self.derived = torch.zeros((4, n), **kw)
        self.pid = torch.zeros((18, n), device=self.device, dtype=torch.float32)
        self.pid_outer = torch.zeros((9, n), device=self.device, dtype=torch.float32)   # heading, energy, balance PIDs
        self.wp_index = torch.zeros((n,), device=self.device, dtype=torch.int32)
        self._mission = None
        self._pending = torch.zeros((4, n), **kw)
        self.c = _cabi.FleetC(n=n, precision=precision, state=self.state.data_ptr(), controls=self.controls.data_ptr(),
                              derived=self.derived.data_ptr(), pid=self.pid.data_ptr())
        self.reset()

    # -- stream plumbing: enqueue on torch's current stream --------------------------------------
    def _bind_stream(self):
        self.ctx.set_stream(torch.cuda.current_stream(self.device).cuda_stream)

    def reset(self, initial_state: Optional[torch.Tensor] = None, mask: Optional[torch.Tensor] = None):
        """Simplified6DOF.reset: initial_state [13][n] or None for level flight at 100 m, 20 m/s."""
        self._bind_stream()
        if initial_state is not None:
            initial_state = as_tensor(initial_state, self.device, self.dtype)
            assert initial_state.shape == (13, self.n)
        if mask is not None:
            mask = as_tensor(mask, self.device, torch.uint8)
        self.ctx.check(self.ctx.lib.acb200_fleet_reset(self.ctx.handle, C.byref(self.c), _ptr(mask), _ptr(initial_state)))
        if mask is None:                       # agent.reset() / mission.start() of the outer levels
            self.pid_outer.zero_(); self.wp_index.zero_()
        else:
            m = mask.bool()
            self.pid_outer[:, m] = 0; self.wp_index[m] = 0
        return self.state

    def set_controls(self, surfaces: torch.Tensor):
        """surfaces [4][n] in ControlSurfaces.to_array order; clipped when the step consumes them."""
        self._pending.copy_(as_tensor(surfaces, self.device, self.dtype))

    def step(self, dt: float):
        self._bind_stream()
        self.ctx.check(self.ctx.lib.acb200_step_surface(self.ctx.handle, C.byref(self.c), _ptr(self._pending), dt,
                                                        self.dt_physics))
        return self.state

    def step_surface(self, surfaces: torch.Tensor, dt: float):
        self.set_controls(surfaces)
        return self.step(dt)

    def _cascade(self, level, cmd, dt, dt_ctrl, surfaces_out):
        self._bind_stream()
        cmd = as_tensor(cmd, self.device, self.dtype)
        assert cmd.shape == (4, self.n)
        self.ctx.check(self.ctx.lib.acb200_step_cascade(self.ctx.handle, C.byref(self.c), level, _ptr(cmd),
                                                        dt if dt_ctrl is None else dt_ctrl, dt, self.dt_physics,
                                                        _ptr(surfaces_out)))
        return self.state

    def step_rate(self, cmd: torch.Tensor, dt: float, dt_ctrl: Optional[float] = None, surfaces_out=None):
        """RateAgent.compute_action(cmd, state, dt_ctrl) + set_controls + step(dt); cmd = p,q,r,throttle."""
        return self._cascade(_cabi.LEVEL_RATE, cmd, dt, dt_ctrl, surfaces_out)

    def step_attitude(self, cmd: torch.Tensor, dt: float, dt_ctrl: Optional[float] = None, surfaces_out=None):
        """AttitudeAgent cascade; cmd = roll, pitch, yaw (NaN = None, +inf = hold current yaw), throttle."""
        return self._cascade(_cabi.LEVEL_ATTITUDE, cmd, dt, dt_ctrl, surfaces_out)

    def step_hsa(self, cmd: torch.Tensor, dt: float, dt_ctrl: Optional[float] = None, surfaces_out=None):
        """HSAAgent.compute_action (hsa_agent.py:119-231) + set_controls + step(dt); cmd [3][n] = heading, speed, altitude.
        The agent's airspeed / altitude / heading are read from ``self.derived`` (kept current by every step and by
        ``reset``): after writing ``self.state`` by hand, call ``reset(init_state=...)`` rather than stepping on."""
        self._bind_stream()
        cmd = as_tensor(cmd, self.device, self.dtype)
        assert cmd.shape == (3, self.n)
        self.ctx.check(self.ctx.lib.acb200_step_hsa(self.ctx.handle, C.byref(self.c), _ptr(self.pid_outer), _ptr(cmd),
                                                    dt if dt_ctrl is None else dt_ctrl, dt, self.dt_physics, _ptr(surfaces_out)))
        return self.state

    def set_mission(self, waypoints, acceptance_radius: float = 40.0, guidance: str = "PP", loop: bool = False, stats: bool = True):
        """MissionPlanner(waypoints, acceptance_radius).start() for every aircraft (mission_planner.py:56-93).
        waypoints [n_wp][4] = north, east, down, speed (NaN = None); guidance 'LOS' | 'PP' (waypoint_agent.py:33-52).
        stats: accumulate the fleet's mission statistics on the device (``mission_stats()``, 16 doubles)."""
        wp = as_tensor(waypoints, self.device, self.dtype)
        assert wp.ndim == 2 and wp.shape[1] == 4 and wp.shape[0] >= 1
        g = {"LOS": _cabi.GUIDANCE_LOS, "PP": _cabi.GUIDANCE_PP, "PURE_PURSUIT": _cabi.GUIDANCE_PP}[guidance.upper()]
        self.waypoints = wp
        self.wp_index.zero_()
        self.mission_stats_buf = torch.zeros((16,), device=self.device, dtype=torch.float64) if stats else None
        self._mission = _cabi.MissionC(waypoints=wp.data_ptr(), n_waypoints=wp.shape[0], guidance=g,
                                       wp_index=self.wp_index.data_ptr(), loop=int(loop), acceptance_radius=acceptance_radius,
                                       stats=None if not stats else self.mission_stats_buf.data_ptr())

    def mission_stats(self, reset: bool = False, allreduce: bool = False) -> dict:
        """The fleet's mission statistics since the last reset of the accumulator: counts and sums (``_cabi.MISSION_STATS_NAMES``)
        plus the derived means.  allreduce: summed over the ranks of the default process group first (the one collective of a
        sharded mission fleet, 128 bytes per report interval)."""
        if self.mission_stats_buf is None:
            raise RuntimeError("set_mission(..., stats=True) first")
        from .sharding import allreduce_episode_stats
        v = (allreduce_episode_stats(self.mission_stats_buf) if allreduce else self.mission_stats_buf).cpu().tolist()
        if reset:
            self.mission_stats_buf.zero_()
        out = dict(zip(_cabi.MISSION_STATS_NAMES, v))
        steps = max(out["aircraft_steps"], 1.0)
        out.update(cross_track_mean=out["cross_track_abs_sum"] / steps, cross_track_rms=(out["cross_track_sq_sum"] / steps) ** 0.5,
                   altitude_error_mean=out["altitude_error_abs_sum"] / steps, airspeed_mean=out["airspeed_sum"] / steps,
                   distance_to_waypoint_mean=out["distance_to_waypoint_sum"] / steps)
        return out

    def step_waypoint(self, dt: float, dt_ctrl: Optional[float] = None, surfaces_out=None, hsa_cmd_out=None):
        """One pass of the loop of examples/03_waypoint_square_demo.py:147-209: MissionPlanner.update,
        WaypointAgent.compute_action, set_controls, step(dt).  Aircraft whose mission is complete are left untouched."""
        if self._mission is None:
            raise RuntimeError("set_mission() first")
        self._bind_stream()
        self.ctx.check(self.ctx.lib.acb200_step_waypoint(self.ctx.handle, C.byref(self.c), _ptr(self.pid_outer),
                                                         C.byref(self._mission), dt if dt_ctrl is None else dt_ctrl, dt,
                                                         self.dt_physics, _ptr(surfaces_out), _ptr(hsa_cmd_out)))
        return self.state

    def mission_complete(self) -> torch.Tensor:
        return self.wp_index >= self._mission.n_waypoints

    def get_state(self):
        return self.state

    def get_backend_type(self) -> str:
        return "simulation"

    def get_dt_nominal(self) -> float:
        return 0.01

    def get_info(self) -> dict:
        return {"backend_type": "simulation", "dt_nominal": 0.01, "physics_engine": "simplified_6dof",
                "dt_physics": self.dt_physics, "aircraft_mass": self.aircraft_params.mass,
                "max_thrust": self.aircraft_params.max_thrust, "n_aircraft": self.n, "precision": self.precision}
