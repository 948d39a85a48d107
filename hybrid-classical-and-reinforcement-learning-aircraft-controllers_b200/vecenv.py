"""SB3-``VecEnv``-shaped front end of ``BatchedRateControlEnv``: what
``make_vec_env`` / ``SubprocVecEnv([make_env(...)] * n)`` give the reference's trainers
(learned_controllers/utils/training_utils.py:23-69, train_overnight.py:136-141), as ONE object over n aircraft.

Protocol mirrored (stable-baselines3 is not installed here; this follows its documented VecEnv contract):
``num_envs``, ``observation_space`` / ``action_space`` (rate_env.py:110-138), ``reset() -> obs``,
``step_async(actions)`` + ``step_wait() -> (obs, rewards, dones, infos)``, ``step(actions)``, ``close()``,
``get_attr / set_attr / env_method``; on ``done`` the returned observation is the RESET observation, the last one of
the episode is ``infos[i]["terminal_observation"]``, ``infos[i]["TimeLimit.truncated"] = truncated and not
terminated``, and (``Monitor``) ``infos[i]["episode"] = {"r", "l", "t"}``.

Every array handed out is numpy over pinned host memory filled by ``acb200_env_step_host``; the per-aircraft
``infos`` list is materialised lazily (a million dicts per step is the reference's cost model, not this one's).
"""
from __future__ import annotations

import time
from typing import Optional

import numpy as np
import torch

from .envs import BatchedRateControlEnv, COMPONENT_NAMES


class Box:
    """The four attributes of gymnasium.spaces.Box that SB3 and the reference read (rate_env.py:110-138)."""

    def __init__(self, low, high, shape=None, dtype=np.float32):
        self.dtype = np.dtype(dtype)
        self.low = np.broadcast_to(np.asarray(low, self.dtype), shape if shape is not None else np.shape(low)).copy()
        self.high = np.broadcast_to(np.asarray(high, self.dtype), self.low.shape).copy()
        self.shape = self.low.shape

    def sample(self, rng=None):
        rng = rng or np.random.default_rng()
        return rng.uniform(self.low, self.high).astype(self.dtype)

    def contains(self, x):
        x = np.asarray(x)
        return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))

    def __repr__(self):
        return f"Box({self.low.min()}, {self.high.max()}, {self.shape}, {self.dtype})"


def observation_space() -> Box:
    """rate_env.py:110-126: 18 float32, bounds as the reference declares them (observations are NOT clipped to them)."""
    hi = np.array([10.0] * 3 + [10.0] * 3 + [20.0] * 3 + [50.0, 500.0] + [np.pi, np.pi / 2, np.pi] + [1.0] * 4, np.float32)
    lo = -hi.copy(); lo[9], lo[10], lo[17] = 5.0, 0.0, 0.0
    return Box(lo, hi)


def action_space() -> Box:
    """rate_env.py:128-138: aileron, elevator, rudder in [-1, 1], throttle in [0, 1]"""
    return Box(np.array([-1, -1, -1, 0], np.float32), np.array([1, 1, 1, 1], np.float32))


class LazyInfos:
    """Sequence of per-aircraft info dicts, built on access."""

    def __init__(self, n, make):
        self._n, self._make = n, make

    def __len__(self):
        return self._n

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self._make(j) for j in range(*i.indices(self._n))]
        if i < 0:
            i += self._n
        if not 0 <= i < self._n:
            raise IndexError(i)
        return self._make(i)

    def __iter__(self):
        return (self._make(i) for i in range(self._n))


class B200VecEnv:
    def __init__(self, num_envs: int, difficulty="medium", command_type="step", episode_length: float = 10.0, dt: float = 0.02,
                 controller: str = "external", precision: int = 32, device=0, seed: int = 0, id_offset: int = 0,
                 lazy_infos: Optional[bool] = None, reward_components: bool = False):
        self.env = BatchedRateControlEnv(num_envs, difficulty=difficulty, episode_length=episode_length, dt=dt,
                                         command_type=command_type, controller=controller, precision=precision, device=device,
                                         seed=seed, id_offset=id_offset, auto_reset=True, track_components=reward_components)
        n = self.num_envs = int(num_envs)
        self.observation_space, self.action_space = observation_space(), action_space()
        self.lazy_infos = (n > 4096) if lazy_infos is None else lazy_infos
        pin = lambda shape, dt_: torch.empty(shape, dtype=dt_, pin_memory=True)
        self._act = pin((n, 4), torch.float32)
        self._obs, self._rew = pin((n, 18), torch.float32), pin((n,), torch.float32)
        self._term, self._trunc = pin((n,), torch.uint8), pin((n,), torch.uint8)
        self._ep_ret, self._ep_len = np.zeros(n, np.float64), np.zeros(n, np.int64)
        self._t0 = time.time()
        self._pending = False
        self._needs_action = controller in ("external", "residual")

    # ---- VecEnv protocol -----------------------------------------------------------------------------------------
    def reset(self):
        self.env.reset()
        self._ep_ret[:] = 0.0; self._ep_len[:] = 0
        self._obs.copy_(self.env.obs)          # synchronous D2H of the reset observations
        return self._obs.numpy().copy()

    def step_async(self, actions) -> None:
        if self._needs_action:
            a = np.asarray(actions, np.float32)
            if a.shape != (self.num_envs, 4):
                raise ValueError(f"actions must have shape ({self.num_envs}, 4), got {a.shape}")
            self._act.numpy()[...] = a
        self._pending = True

    def step_wait(self):
        assert self._pending, "step_wait() without step_async()"
        self._pending = False
        self.env.step_host(self._act if self._needs_action else None, self._obs, self._rew, self._term, self._trunc)
        obs, rew = self._obs.numpy(), self._rew.numpy()
        term, trunc = self._term.numpy().astype(bool), self._trunc.numpy().astype(bool)
        dones = term | trunc
        self._ep_ret += rew; self._ep_len += 1
        finished = np.flatnonzero(dones)
        ep_r, ep_l = self._ep_ret[finished].copy(), self._ep_len[finished].copy()
        self._ep_ret[finished] = 0.0; self._ep_len[finished] = 0
        term_obs = self.env.terminal_obs[torch.as_tensor(finished, device=self.env.device)].cpu().numpy() if finished.size else None
        comps = self.env.components.float().cpu().numpy() if self.env.components is not None else None
        lookup = {int(i): k for k, i in enumerate(finished)}
        elapsed = time.time() - self._t0

        # The reference's per-step info keys (rate_env.py:410-435): time, step, position, rate_command, rate_error, airspeed,
        # altitude, is_settled.  Pulled from the device ONCE per step and only if some info dict is actually built.  For an
        # aircraft that finished this step the device already holds the NEW episode (SB3 auto-reset), so its values are
        # rebuilt from what was kept of the old one: the terminal observation (float32) and the episode length; the
        # terminal position is not kept anywhere and reads NaN.
        env, dt = self.env, self.env.dt
        cache = {}

        def extras():
            if not cache:
                st = env.state.double().cpu().numpy()
                cache.update(time=env.env_time.double().cpu().numpy(), step=env.step_count.cpu().numpy(), pos=st[0:3],
                             cmd=env.rate_cmd.double().cpu().numpy(), rates=st[9:12], der=env.derived.double().cpu().numpy(),
                             settled=env.is_settled.cpu().numpy())
            return cache

        def make(i):
            x = extras()
            k = lookup.get(i)
            if k is None:
                info = {"time": float(x["time"][i]), "step": int(x["step"][i]), "position": x["pos"][:, i].copy(),
                        "rate_command": x["cmd"][:, i].copy(), "rate_error": x["cmd"][:, i] - x["rates"][:, i],
                        "airspeed": float(x["der"][0, i]), "altitude": float(x["der"][1, i]), "is_settled": bool(x["settled"][i])}
            else:
                to = term_obs[k].astype(np.float64)
                info = {"time": float(ep_l[k] * dt), "step": int(ep_l[k]), "position": np.full(3, np.nan),
                        "rate_command": to[3:6].copy(), "rate_error": to[6:9].copy(), "airspeed": float(to[9]),
                        "altitude": float(to[10]), "is_settled": False}
            if comps is not None:     # info["reward_components"], rate_env.py:410-435 (read by the flight plugin's callback)
                info["reward_components"] = {nm: float(comps[k_, i]) for k_, nm in enumerate(COMPONENT_NAMES)}
            if k is not None:
                info["terminal_observation"] = term_obs[k]
                info["TimeLimit.truncated"] = bool(trunc[i] and not term[i])
                info["episode"] = {"r": float(ep_r[k]), "l": int(ep_l[k]), "t": round(elapsed, 6)}
            return info

        infos = LazyInfos(self.num_envs, make) if self.lazy_infos else [make(i) for i in range(self.num_envs)]
        return obs.copy(), rew.copy(), dones, infos

    def step(self, actions):
        self.step_async(actions)
        return self.step_wait()

    def close(self) -> None:
        self.env.ctx.close()

    # ---- attribute plumbing (SB3's get_attr / set_attr / env_method) ------------------------------------------------
    def get_attr(self, name: str, indices=None):
        idx = range(self.num_envs) if indices is None else ([indices] if isinstance(indices, int) else indices)
        if name == "difficulty_scale":
            v = self.env.difficulty.cpu().numpy()
            return [float(v[i]) for i in idx]
        if name == "rate_command":
            v = self.env.rate_cmd.double().cpu().numpy()
            return [v[:, i].copy() for i in idx]
        val = getattr(self.env, name)
        return [val for _ in idx]

    def set_attr(self, name: str, value, indices=None) -> None:
        """``difficulty`` / ``command_type``: the curriculum switch that the reference does by re-creating its envs
        (train_rate.py:150-180)."""
        if name in ("difficulty", "command_type"):
            mask = None
            if indices is not None:
                mask = torch.zeros(self.num_envs, dtype=torch.bool)
                mask[[indices] if isinstance(indices, int) else list(indices)] = True
            cur_d, cur_c = self.env.difficulty.clone(), self.env.command_type.clone()
            self.env.set_curriculum(value if name == "difficulty" else cur_d, value if name == "command_type" else cur_c, mask)
        else:
            setattr(self.env, name, value)

    def env_method(self, method_name: str, *args, indices=None, **kwargs):
        out = getattr(self.env, method_name)(*args, **kwargs)
        n = self.num_envs if indices is None else (1 if isinstance(indices, int) else len(indices))
        return [out] * n

    def seed(self, seed: Optional[int] = None):
        if seed is not None:
            self.env.cfg.seed = int(seed)
        return [self.env.cfg.seed + i for i in range(self.num_envs)]

    def episode_stats(self, reset: bool = False) -> dict:
        return self.env.episode_stats(reset)
