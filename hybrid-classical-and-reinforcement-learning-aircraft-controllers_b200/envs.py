"""Batched rate-tracking environment: N RateControlEnv instances (learned_controllers/envs/rate_env.py
in the reference) advanced by one fused kernel per control step, with SB3-VecEnv auto-reset semantics
and the curriculum (difficulty, command type) as per-aircraft tensors."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _cabi
from .fleet import _dtype, _ptr, as_tensor

DIFFICULTY_SCALE = {"easy": 0.3, "medium": 0.5, "hard": 0.7}          # generators.py:41-45
COMMAND_TYPES = {"step": _cabi.CMD_STEP, "ramp": _cabi.CMD_RAMP, "sine": _cabi.CMD_SINE, "random": _cabi.CMD_RANDOM}
CONTROLLERS = {"external": _cabi.CTRL_EXTERNAL, "rate_pid": _cabi.CTRL_RATE_PID, "attitude_pid": _cabi.CTRL_ATTITUDE_PID,
               "residual": _cabi.CTRL_RESIDUAL}
COMPONENT_NAMES = ("tracking", "smoothness", "stability", "oscillation", "survival", "settle_bonus", "crash_penalty",
                   "residual_bonus")
STATS_NAMES = ("episodes", "return_sum", "return_sq_sum", "length_sum", "crashes", "truncations") + tuple(
    "sum_" + c for c in COMPONENT_NAMES) + ("nonfinite_repairs", "aircraft_steps")


RANDOMIZED_FIELDS = ("mass", "inertia_xx", "inertia_yy", "inertia_zz", "cl_alpha", "cd_0", "cl_elevator", "cm_elevator",
                     "cn_rudder", "cl_aileron", "cm_alpha", "cn_beta", "cl_beta", "damping_roll", "damping_pitch", "damping_yaw",
                     "max_thrust")


def randomize_aircraft_params(base, n_groups: int, spread: float = 0.2, seed: int = 0, fields=RANDOMIZED_FIELDS):
    """``n_groups`` copies of ``base`` with each listed field scaled by U(1 - spread, 1 + spread) (numpy PCG64, host side:
    a few hundred numbers).  Group 0 is ``base`` itself."""
    import numpy as np
    rng = np.random.default_rng(seed)
    out = []
    for g in range(n_groups):
        p = _cabi.AircraftParamsC.from_buffer_copy(bytes(base))
        if g:
            for f in fields:
                setattr(p, f, getattr(base, f) * float(rng.uniform(1.0 - spread, 1.0 + spread)))
        out.append(p)
    return out


def randomize_aircraft_params_table(base, n: int, spread: float = 0.2, seed: int = 0, fields=RANDOMIZED_FIELDS, id_offset: int = 0):
    """One ``AircraftParams`` per aircraft: a ctypes array of n ``AircraftParamsC`` with each listed field of ``base`` scaled by
    U(1 - spread, 1 + spread), drawn per GLOBAL aircraft id (a hash of (seed, id, field), so the draw does not depend on
    the sharding).  Vectorised on the host: a million aircraft are a few hundred milliseconds."""
    import numpy as np
    nf = len(_cabi.AIRCRAFT_FIELDS)
    tab = np.tile(np.frombuffer(bytes(base), dtype=np.float64), (n, 1))
    ids = np.arange(id_offset, id_offset + n, dtype=np.uint64)
    with np.errstate(over="ignore"):
        for f in fields:                         # splitmix64 of (seed, global id, field): shard-independent uniforms
            k = _cabi.AIRCRAFT_FIELDS.index(f)
            x = ids * np.uint64(0x9E3779B97F4A7C15) + np.uint64((seed * 0xD1B54A32D192ED03 + (k + 1) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF)
            x ^= x >> np.uint64(30); x *= np.uint64(0xBF58476D1CE4E5B9)
            x ^= x >> np.uint64(27); x *= np.uint64(0x94D049BB133111EB)
            x ^= x >> np.uint64(31)
            u = (x >> np.uint64(11)).astype(np.float64) / float(1 << 53)
            tab[:, k] *= 1.0 - spread + 2.0 * spread * u
    assert C.sizeof(_cabi.AircraftParamsC) == nf * 8
    return (_cabi.AircraftParamsC * n).from_buffer_copy(np.ascontiguousarray(tab).tobytes())


class BatchedRateControlEnv:
    """obs float32 [n,18]; action float32 [n,4] = aileron, elevator, rudder, throttle (rate_env.py:228-233)."""

    def __init__(self, n: int, difficulty="medium", episode_length: float = 10.0, dt: float = 0.02,
                 command_type="step", controller: str = "external", precision: int = 32, device=0, seed: int = 0,
                 id_offset: int = 0, auto_reset: bool = True, pid_throttle: float = 0.6, residual_scale: float = 0.3,
                 track_components: bool = False, ctx: Optional[_cabi.Context] = None):
        if not torch.cuda.is_available():
            raise RuntimeError("acb200 needs a CUDA device (no CPU fallback)")
        self.device = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
        self.n, self.num_envs, self.precision, self.dtype = int(n), int(n), precision, _dtype(precision)
        self.dt, self.episode_length = dt, episode_length
        self.ctx = ctx or _cabi.Context(self.device.index or 0)
        self.cfg = _cabi.EnvConfigC(dt=dt, dt_physics=0.001, max_steps=int(episode_length / dt),
                                    controller=CONTROLLERS[controller], auto_reset=int(auto_reset),
                                    pid_throttle=pid_throttle, residual_scale=residual_scale, seed=seed, id_offset=id_offset)
        d, r = self.device, self.dtype
        z = lambda shape, dt_: torch.zeros(shape, device=d, dtype=dt_)
        self.state, self.controls, self.derived = z((13, n), r), z((4, n), r), z((4, n), r)
        self.pid = z((18, n), torch.float32)
        self.att_cmd, self.rate_cmd = z((4, n), r), z((3, n), r)
        self.prev_action = z((4, n), torch.float32)
        self.rew_mem, self.sched, self.env_time = z((7, n), r), z((7, n), r), z((n,), r)
        self.step_count = z((n,), torch.int32)
        self.is_settled, self.command_type = z((n,), torch.uint8), z((n,), torch.uint8)
        self.difficulty = z((n,), torch.float32)
        self.episode_idx = z((n,), torch.int32)          # uint32 on the device; torch has no uint32 arithmetic
        self.ep_return = z((n,), r)
        self.obs = z((n, 18), torch.float32)
        self.reward = z((n,), r)
        self.terminated, self.truncated = z((n,), torch.uint8), z((n,), torch.uint8)
        self.terminal_obs = z((n, 18), torch.float32)
        self.components = z((8, n), r) if track_components else None
        self.stats = z((16,), torch.float64)
        self._scratch = z((n * 5,), torch.float32)       # staged action + float reward for step_host
        self.set_curriculum(difficulty, command_type)
        self.c = _cabi.EnvC()
        self.c.fleet = _cabi.FleetC(n=n, precision=precision, state=self.state.data_ptr(), controls=self.controls.data_ptr(),
                                    derived=self.derived.data_ptr(), pid=self.pid.data_ptr())
        for name in _cabi.ENV_POINTER_FIELDS:
            t = getattr(self, name)
            setattr(self.c, name, None if t is None else t.data_ptr())

    # -- curriculum: a tensor write, not an env re-creation (train_rate.py:150-180 re-creates envs) ----
    def set_curriculum(self, difficulty, command_type, mask: Optional[torch.Tensor] = None):
        dv = difficulty if not isinstance(difficulty, str) else DIFFICULTY_SCALE[difficulty]
        cv = command_type if not isinstance(command_type, str) else COMMAND_TYPES[command_type]
        dv = torch.as_tensor(dv, device=self.device, dtype=torch.float32).expand(self.n)
        cv = torch.as_tensor(cv, device=self.device, dtype=torch.uint8).expand(self.n)
        if mask is None:
            self.difficulty.copy_(dv); self.command_type.copy_(cv)
        else:
            m = as_tensor(mask, self.device).bool()
            self.difficulty[m] = dv[m]; self.command_type[m] = cv[m]

    def _bind_stream(self):
        self.ctx.set_stream(torch.cuda.current_stream(self.device).cuda_stream)

    def reset(self, mask: Optional[torch.Tensor] = None):
        self._bind_stream()
        if mask is not None:
            mask = as_tensor(mask, self.device, torch.uint8)
        self.ctx.check(self.ctx.lib.acb200_env_reset(self.ctx.handle, C.byref(self.c), C.byref(self.cfg), _ptr(mask)))
        return self.obs

    # -- domain randomisation: parameter GROUPS ----------------------------
    def set_param_groups(self, groups, bounds=None):
        """Give contiguous groups of aircraft their own ``AircraftParams`` (the reference has one per env instance,
        simulation_backend.py:55-80).  ``groups``: list of ``_cabi.AircraftParamsC``; ``bounds``: len(groups)+1 offsets,
        multiples of 128 except the last (default: an even split).  Each group is one launch with its parameters in the
        constant bank, on EVERY step path (``step``, ``step_host``, the VecEnv shim): the groups live in the context
        (``acb200_set_param_groups``), which validates all of them first -- a bad group raises ``ValueError`` like
        ``AircraftParams.__post_init__`` and changes nothing.  ``None`` / [] returns to the context's one parameter set."""
        lib, h = self.ctx.lib, self.ctx.handle
        if not groups:
            self.ctx.check(lib.acb200_set_param_groups(h, None, None, 0))
            self._groups = None
            return
        g = len(groups)
        if bounds is None:
            per = ((self.n + g - 1) // g + 127) // 128 * 128
            bounds = [min(k * per, self.n) for k in range(g)] + [self.n]
        bounds = [int(b) for b in bounds]
        if len(bounds) != g + 1 or bounds[-1] != self.n:
            raise ValueError("bounds must rise from 0 to n in multiples of 128")
        arr = (_cabi.AircraftParamsC * g)(*groups)
        b = (C.c_int64 * (g + 1))(*bounds)
        self.ctx.check(lib.acb200_set_param_groups(h, arr, b, g))
        self._groups = (list(groups), bounds)

    def set_param_table(self, params):
        """Per-AIRCRAFT ``AircraftParams``: ``params`` is a sequence / ctypes array of n
        ``_cabi.AircraftParamsC`` (``randomize_aircraft_params_table``), or None to return to one parameter set.  The
        folded table ([words][n] of the build's real type) lives in a torch tensor this env owns; every step path then runs the
        table variant of the kernel.  A bad row raises ``ValueError`` like ``AircraftParams.__post_init__`` and changes nothing."""
        lib, h = self.ctx.lib, self.ctx.handle
        if params is None:
            self.ctx.check(lib.acb200_set_param_table(h, None, 0, 0))
            self.param_table = None
            return
        if len(params) != self.n:
            raise ValueError(f"need one AircraftParams per aircraft: {len(params)} given for {self.n}")
        arr = params if isinstance(params, C.Array) else (_cabi.AircraftParamsC * self.n)(*params)
        table = torch.zeros((lib.acb200_param_table_words(), self.n), device=self.device, dtype=self.dtype)
        self._bind_stream()
        self.ctx.check(lib.acb200_build_param_table(h, arr, self.n, 0, self.n, self.precision, _ptr(table)))
        self.ctx.check(lib.acb200_set_param_table(h, _ptr(table), self.n, self.precision))
        self.param_table = table

    def step(self, action: Optional[torch.Tensor] = None):
        """One control step for every aircraft; tensors returned are views of the env's own buffers."""
        self._bind_stream()
        if action is not None:
            action = as_tensor(action, self.device)
            assert action.dtype == torch.float32 and action.shape == (self.n, 4) and action.is_contiguous()
        self.ctx.check(self.ctx.lib.acb200_env_step(self.ctx.handle, C.byref(self.c), C.byref(self.cfg), _ptr(action)))
        return self.obs, self.reward, self.terminated, self.truncated

    def step_host(self, action_host, obs_host, reward_host, terminated_host, truncated_host, att_cmd_host=None):
        """Same step with (pinned) HOST tensors in and out; synchronous.  att_cmd_host [4][n] (dtype of the build)
        is this step's attitude command for controller="attitude_pid".  Any OUTPUT may be None: it is then not copied
        (reward + flags alone are 6 B per aircraft instead of 78 with the observation); ``gather_done`` returns the
        finished aircraft's observations."""
        self._bind_stream()
        if att_cmd_host is not None:
            assert att_cmd_host.dtype == self.dtype and att_cmd_host.shape == (4, self.n) and att_cmd_host.is_contiguous()
        self.ctx.check(self.ctx.lib.acb200_env_step_host(
            self.ctx.handle, C.byref(self.c), C.byref(self.cfg), _ptr(action_host), _ptr(att_cmd_host), _ptr(self._scratch),
            _ptr(obs_host), _ptr(reward_host), _ptr(terminated_host), _ptr(truncated_host)))
        return obs_host, reward_host, terminated_host, truncated_host

    def gather_done(self, capacity: Optional[int] = None, to_host: bool = True):
        """The aircraft that finished an episode in the last step, ascending ids: (ids int64 [k], terminal_obs float32
        [k,18], reset_obs float32 [k,18]).  One compaction on the device (``acb200_env_gather_done``) and, with
        ``to_host``, copies of k rows -- not of the fleet."""
        self._bind_stream()
        cap = int(capacity if capacity is not None else max(1024, self.n // 16))
        if getattr(self, "_done_cap", 0) < cap:
            d = self.device
            self._done_ids = torch.empty((cap,), device=d, dtype=torch.int64)
            self._done_term = torch.empty((cap, 18), device=d, dtype=torch.float32)
            self._done_reset = torch.empty((cap, 18), device=d, dtype=torch.float32)
            self._done_count = torch.zeros((1,), device=d, dtype=torch.int64)
            self._done_cap = cap
        self.ctx.check(self.ctx.lib.acb200_env_gather_done(self.ctx.handle, C.byref(self.c), cap, _ptr(self._done_ids),
                                                          _ptr(self._done_term), _ptr(self._done_reset), _ptr(self._done_count)))
        k = int(self._done_count.item())
        if k > cap:                      # rare: more finished than the buffers hold -- grow and gather again
            return self.gather_done(capacity=k, to_host=to_host)
        out = (self._done_ids[:k], self._done_term[:k], self._done_reset[:k])
        return tuple(t.cpu() for t in out) if to_host else out

    # -- checkpoint: everything a resumed run needs to continue bit for bit -------------------
    _STATE_TENSORS = ("state", "controls", "derived", "pid", "att_cmd", "rate_cmd", "prev_action", "rew_mem", "sched", "env_time",
                      "step_count", "is_settled", "command_type", "difficulty", "episode_idx", "ep_return", "obs", "reward",
                      "terminated", "truncated", "terminal_obs", "stats")

    def state_dict(self) -> dict:
        """The SoA tensors (incl. the Philox counter word ``episode_idx``) + the config words the random streams are keyed
        by.  CPU copies: a checkpoint written on one GPU loads on another."""
        sd = {k: getattr(self, k).detach().cpu().clone() for k in self._STATE_TENSORS}
        sd["config"] = {"n": self.n, "precision": self.precision, "dt": self.cfg.dt, "dt_physics": self.cfg.dt_physics,
                        "max_steps": self.cfg.max_steps, "controller": self.cfg.controller, "auto_reset": self.cfg.auto_reset,
                        "pid_throttle": self.cfg.pid_throttle, "residual_scale": self.cfg.residual_scale, "seed": self.cfg.seed,
                        "id_offset": self.cfg.id_offset}
        return sd

    def load_state_dict(self, sd: dict):
        c = sd["config"]
        if c["n"] != self.n or c["precision"] != self.precision:
            raise ValueError(f"checkpoint is for n={c['n']} precision={c['precision']}, env is n={self.n} precision={self.precision}")
        for k in self._STATE_TENSORS:
            getattr(self, k).copy_(sd[k].to(self.device))
        for k in ("dt", "dt_physics", "max_steps", "controller", "auto_reset", "pid_throttle", "residual_scale", "seed", "id_offset"):
            setattr(self.cfg, k, c[k])
        self.dt = self.cfg.dt

    def episode_stats(self, reset: bool = False) -> dict:
        v = self.stats.cpu().tolist()
        if reset:
            self.stats.zero_()
        return dict(zip(STATS_NAMES, v))
