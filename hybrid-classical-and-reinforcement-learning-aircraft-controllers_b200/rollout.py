"""Device-resident rollout storage with GAE on the device: the stable-baselines3
``RolloutBuffer`` the reference's trainers fill from ``SubprocVecEnv`` (learned_controllers/utils/training_utils.py:
49-69,112-160), as CUDA tensors that never visit the host.  Method names and argument meaning follow SB3's class
(``reset``, ``add``, ``compute_returns_and_advantage``, ``get``); the advantage recursion is ``acb200_gae``
(csrc/rollout.cu), bit-identical to SB3's float32 numpy loop."""
from __future__ import annotations

from typing import Iterator, Optional

import torch

from . import _cabi

OBS_DIM, ACT_DIM = 18, 4


class RolloutBufferB200:
    def __init__(self, buffer_size: int, n_envs: int, device=0, gamma: float = 0.99, gae_lambda: float = 0.95,
                 ctx: Optional[_cabi.Context] = None):
        self.buffer_size, self.n_envs = int(buffer_size), int(n_envs)
        self.gamma, self.gae_lambda = float(gamma), float(gae_lambda)
        self.device = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
        self.ctx = ctx or _cabi.Context(self.device.index or 0)
        T, n = self.buffer_size, self.n_envs
        z = lambda *shape, dtype=torch.float32: torch.zeros(shape, dtype=dtype, device=self.device)
        self.observations, self.actions = z(T, n, OBS_DIM), z(T, n, ACT_DIM)
        self.rewards, self.values, self.log_probs = z(T, n), z(T, n), z(T, n)
        self.episode_starts = z(T, n, dtype=torch.uint8)
        self.advantages, self.returns = z(T, n), z(T, n)
        self.pos, self.full, self.generator_ready = 0, False, False

    def reset(self) -> None:
        self.pos, self.full, self.generator_ready = 0, False, False

    def add(self, obs, action, reward, episode_start, value, log_prob=None) -> None:
        """One env step of every environment; all arguments are CUDA tensors (device-to-device copies only)."""
        if self.full:
            raise RuntimeError("rollout buffer is full: compute_returns_and_advantage(), get(), reset()")
        t = self.pos
        self.observations[t].copy_(obs); self.actions[t].copy_(action)
        self.rewards[t].copy_(reward.reshape(-1)); self.values[t].copy_(value.reshape(-1))
        self.episode_starts[t].copy_(episode_start.reshape(-1))
        if log_prob is not None:
            self.log_probs[t].copy_(log_prob.reshape(-1))
        self.pos += 1
        self.full = self.pos == self.buffer_size

    def compute_returns_and_advantage(self, last_values: torch.Tensor, dones: torch.Tensor) -> None:
        lv = last_values.detach().reshape(-1).to(self.device, torch.float32).contiguous()
        d = dones.reshape(-1).to(self.device, torch.uint8).contiguous()
        if lv.numel() != self.n_envs or d.numel() != self.n_envs:
            raise ValueError("last_values and dones must have one entry per environment")
        self.ctx.set_stream(torch.cuda.current_stream(self.device).cuda_stream)
        self.ctx.check(self.ctx.lib.acb200_gae(
            self.ctx.handle, self.buffer_size, self.n_envs, self.rewards.data_ptr(), self.values.data_ptr(),
            self.episode_starts.data_ptr(), lv.data_ptr(), d.data_ptr(), self.gamma, self.gae_lambda,
            self.advantages.data_ptr(), self.returns.data_ptr()))

    def get(self, batch_size: Optional[int] = None, generator: Optional[torch.Generator] = None) -> Iterator[dict]:
        """Shuffled minibatches over the flattened (T*n) samples, as SB3's ``get`` yields them (its swap_and_flatten
        order: environment-major); tensors stay on the device."""
        if not self.full:
            raise RuntimeError("rollout buffer is not full")
        N = self.buffer_size * self.n_envs
        flat = lambda x: x.transpose(0, 1).reshape(N, *x.shape[2:])
        data = {"observations": flat(self.observations), "actions": flat(self.actions), "old_values": flat(self.values),
                "old_log_prob": flat(self.log_probs), "advantages": flat(self.advantages), "returns": flat(self.returns)}
        perm = torch.randperm(N, device=self.device, generator=generator)
        bs = N if batch_size is None else int(batch_size)
        for s in range(0, N, bs):
            idx = perm[s:s + bs]
            yield {k: v[idx] for k, v in data.items()}
