"""``RateControlEnvB200`` / ``ResidualRateControlEnvB200``: the reference's ``gym.Env`` API over a one-aircraft fleet.

Mirrors learned_controllers/envs/rate_env.py:18-476 and residual_rate_env.py:17-172: constructor keywords
(``difficulty, episode_length, dt, command_type, render_mode, rng_seed`` [+ ``residual_scale``]), ``reset(seed, options) ->
(obs, info)``, ``step(action) -> (obs, reward, terminated, truncated, info)``, the attributes callers read (``.sim``,
``.rate_command``, ``.dt``, ``.max_steps``, ``.step_count``, ``.current_time``, ``.settle_bonus.is_settled``,
``.observation_space``, ``.action_space``, ``.episode_rewards``; callers: residual_rate_env.py:112-127,
utils/pid_demonstrations.py:57-60, eval_rate.py:104-108, tensorboard_flight_plugin/.../callbacks.py:102,242) and the info
keys ``time, step, position, rate_command, rate_error, airspeed, altitude, is_settled, reward_components``.

The arithmetic is one launch of ``env_step_kernel`` on a fleet of ONE aircraft per ``step`` (the fp64 parity build by
default) -- this object is for the reference's examples, evaluation scripts and tests; a trainer uses ``B200VecEnv`` /
``BatchedRateControlEnv``, where a step of a million aircraft costs what one step costs here.
``rate_command`` can be assigned (eval_rate.py:104-108 does); ``reset(seed=s)`` re-keys the Philox stream, so equal seeds give
equal episodes (the reference re-seeds only one of its three MT19937 streams there, SURVEY.md Appendix A item 25 -- stream
parity is impossible by design, distribution parity is tested in tests/test_reset_sampler.py).
"""
from __future__ import annotations

from typing import Any, Dict, Optional, Tuple

import numpy as np
import torch

from .adapters import AircraftState
from .envs import BatchedRateControlEnv, COMPONENT_NAMES
from .vecenv import Box, action_space, observation_space

# the keys RateTrackingReward.compute returns (rewards.py:126-137), in the kernel's component order; "total" is their sum
_TRACKER_KEYS = ("tracking", "smoothness", "stability", "oscillation", "survival")


class _SimView:
    """``env.sim``: what callers use of SimulationAircraftBackend on the env's own aircraft -- ``get_state()``
    (residual_rate_env.py:112, pid_demonstrations.py:57) and the info accessors."""

    def __init__(self, env: BatchedRateControlEnv):
        self._env = env

    def get_state(self) -> AircraftState:
        s = self._env.state[:, 0].double().cpu().numpy()
        d = self._env.derived[:, 0].double().cpu().numpy()
        return AircraftState(time=float(s[12]), position=s[0:3].copy(), velocity=s[3:6].copy(), attitude=s[6:9].copy(),
                             angular_rate=s[9:12].copy(), airspeed=float(d[0]), altitude=float(d[1]), ground_speed=float(d[2]),
                             heading=float(d[3]))

    def get_backend_type(self) -> str:
        return "simulation"

    def get_dt_nominal(self) -> float:
        return 0.01

    def get_info(self) -> dict:
        return {"backend_type": "simulation", "dt_nominal": 0.01, "physics_engine": "simplified_6dof",
                "dt_physics": self._env.cfg.dt_physics, "aircraft_mass": 8.0, "max_thrust": 50.0}


class _SettleView:
    """``env.settle_bonus``: ``is_settled`` / ``settle_timer`` of SettlingTimeBonus (rewards.py:187-188) as the kernel holds them."""

    def __init__(self, env: BatchedRateControlEnv):
        self._env = env

    @property
    def is_settled(self) -> bool:
        return bool(self._env.is_settled[0].item())

    @property
    def settle_timer(self) -> float:
        return float(self._env.rew_mem[6, 0].item())


class RateControlEnvB200:
    """``RateControlEnv`` (rate_env.py:18-476), N = 1."""

    metadata = {"render_modes": ["human"], "render_fps": 50}
    _controller = "external"

    def __init__(self, difficulty: str = "medium", episode_length: float = 10.0, dt: float = 0.02, command_type: str = "step",
                 render_mode: Optional[str] = None, rng_seed: Optional[int] = None, precision: int = 64, device=0, **batched_kw):
        if difficulty not in ("easy", "medium", "hard"):
            raise KeyError(difficulty)                                   # the reference's dict lookup, generators.py:41-45
        self.difficulty, self.episode_length, self.dt = difficulty, episode_length, dt
        # an unknown command type falls back to a zero command in the reference (rate_env.py:338-340); a constant step
        # command of magnitude zero is not what "step" draws, so refuse instead of guessing
        if command_type not in ("step", "ramp", "sine", "random"):
            raise ValueError(f"unknown command_type {command_type!r}")
        self.command_type, self.render_mode = command_type, render_mode
        self._env = BatchedRateControlEnv(1, difficulty=difficulty, episode_length=episode_length, dt=dt, command_type=command_type,
                                          controller=self._controller, precision=precision, device=device,
                                          seed=0 if rng_seed is None else int(rng_seed), auto_reset=False, track_components=True,
                                          **batched_kw)
        self.max_steps = int(episode_length / dt)
        self.sim = _SimView(self._env)
        self.settle_bonus = _SettleView(self._env)
        self.observation_space, self.action_space = observation_space(), action_space()
        self._max_rates = np.radians([180.0, 180.0, 160.0])
        self._action = torch.zeros((1, 4), device=self._env.device, dtype=torch.float32)
        self.prev_action = np.zeros(4)
        self.episode_rewards = {k: 0.0 for k in ("tracking", "smoothness", "stability", "oscillation", "survival", "settle_bonus",
                                                 "crash_penalty")}

    # ---- attributes the reference's callers read (and eval scripts assign) -------------------------------------------
    @property
    def rate_command(self) -> np.ndarray:
        return self._env.rate_cmd[:, 0].double().cpu().numpy()

    @rate_command.setter
    def rate_command(self, value) -> None:
        v = np.asarray(value, np.float64).reshape(3)
        self._env.rate_cmd[:, 0] = torch.as_tensor(v, device=self._env.device, dtype=self._env.dtype)

    @property
    def current_time(self) -> float:
        return float(self._env.env_time[0].item())

    @property
    def step_count(self) -> int:
        return int(self._env.step_count[0].item())

    @property
    def batched(self) -> BatchedRateControlEnv:
        """the one-aircraft fleet underneath (device tensors, C-ABI descriptors)"""
        return self._env

    # ---- gym.Env ----------------------------------------------------------------------------------------------------
    def _info(self, components: Optional[dict] = None) -> Dict[str, Any]:
        e = self._env
        s = e.state[:, 0].double().cpu().numpy()
        d = e.derived[:, 0].double().cpu().numpy()
        cmd = self.rate_command
        info = {"time": self.current_time, "step": self.step_count, "position": s[0:3].copy(), "rate_command": cmd,
                "rate_error": cmd - s[9:12], "airspeed": float(d[0]), "altitude": float(d[1]),
                "is_settled": self.settle_bonus.is_settled}
        if components is not None:
            info["reward_components"] = components
        return info

    def reset(self, seed: Optional[int] = None, options: Optional[Dict[str, Any]] = None) -> Tuple[np.ndarray, Dict[str, Any]]:
        if seed is not None:                       # equal seeds -> equal episodes: re-key the stream and rewind its counter
            self._env.cfg.seed = int(seed)
            self._env.episode_idx.zero_()
        self._env.reset()
        self.prev_action = np.array([0.0, 0.0, 0.0, 0.5])
        self.episode_rewards = {k: 0.0 for k in self.episode_rewards}
        return self._env.obs[0].cpu().numpy().copy(), self._info()

    def _step_device(self, action) -> Tuple[np.ndarray, float, bool, bool, dict]:
        a = np.asarray(action, np.float32).reshape(4)
        self._action.copy_(torch.as_tensor(a).reshape(1, 4))
        obs, rew, term, trunc = self._env.step(self._action)
        comp = self._env.components[:, 0].double().cpu().numpy()
        terminated, truncated = bool(term[0].item()), bool(trunc[0].item())
        components = {k: float(comp[i]) for i, k in enumerate(_TRACKER_KEYS)}
        components["total"] = float(sum(components[k] for k in _TRACKER_KEYS))       # rewards.py:123
        components["tracking_error_mse"] = -2.0 * components["tracking"]             # r_tracking = -0.5 * mse, rewards.py:76-79
        if terminated and not truncated:
            components["crash_penalty"] = float(comp[6])                             # rate_env.py:289-294
        for k in _TRACKER_KEYS:
            self.episode_rewards[k] += components[k]
        self.episode_rewards["settle_bonus"] += float(comp[5])
        self.prev_action = self._env.prev_action[:, 0].cpu().numpy().astype(np.float32)
        return obs[0].cpu().numpy().copy(), float(rew[0].item()), terminated, truncated, components

    def step(self, action) -> Tuple[np.ndarray, float, bool, bool, Dict[str, Any]]:
        obs, reward, terminated, truncated, components = self._step_device(action)
        return obs, reward, terminated, truncated, self._info(components)

    def render(self):
        return None

    def close(self):
        pass


class ResidualRateControlEnvB200(RateControlEnvB200):
    """``ResidualRateControlEnv`` (residual_rate_env.py:17-172): the action is a residual in [-1, 1]^4, scaled by
    ``residual_scale`` and added to the rate PID's output inside the kernel (controller mode RESIDUAL, csrc/env.cuh
    ``env_action``), which also adds the small-residual bonus to the reward."""

    _controller = "residual"

    def __init__(self, difficulty: str = "medium", episode_length: float = 10.0, dt: float = 0.02, command_type: str = "step",
                 render_mode: Optional[str] = None, rng_seed: Optional[int] = None, residual_scale: float = 0.3, precision: int = 64,
                 device=0):
        super().__init__(difficulty=difficulty, episode_length=episode_length, dt=dt, command_type=command_type,
                         render_mode=render_mode, rng_seed=rng_seed, precision=precision, device=device,
                         residual_scale=residual_scale)
        self.residual_scale = residual_scale
        self.base_env = self                      # callers reach through .base_env.sim / .base_env.rate_command / .base_env.dt
        self.action_space = Box(np.array([-1, -1, -1, -1], np.float32), np.array([1, 1, 1, 1], np.float32))   # :74-78
        self.last_pid_action = np.zeros(4)

    def reset(self, seed: Optional[int] = None, options: Optional[Dict[str, Any]] = None):
        obs, info = super().reset(seed=seed, options=options)
        self.last_pid_action = np.zeros(4)
        info["pid_action"] = self.last_pid_action.copy()                              # :92-93
        return obs, info

    def step(self, residual_action):
        res = np.asarray(residual_action, np.float32).reshape(4)
        obs, reward, terminated, truncated, components = self._step_device(res)
        info = self._info(components)
        # :148-150.  The kernel keeps the clipped combined action (the env's prev_action); the PID's own share is what is
        # left after taking the scaled residual off wherever the clip did not bind.
        scaled = res * np.float32(self.residual_scale)
        combined = self.prev_action.copy()
        info["residual_action"] = scaled
        info["combined_action"] = combined
        info["pid_action"] = self.last_pid_action = (combined - scaled).astype(np.float32)
        return obs, reward, terminated, truncated, info
