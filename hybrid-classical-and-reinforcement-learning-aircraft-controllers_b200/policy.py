"""Learned rate policy, actor forward on the GPU: the host-side mirror of what the reference gets
from ``model.predict(obs, state=lstm_state, episode_start=..., deterministic=True)``
(controllers/learned_rate_agent.py:220-227) for the two architectures of learned_controllers/networks/lstm_policy.py.

Weights come from a torch ``state_dict`` with stable-baselines3 / sb3-contrib parameter names
(``features_extractor.embedding.0.weight``, ``features_extractor.lstm.weight_ih_l0``, ``lstm_actor.weight_hh_l0``,
``mlp_extractor.policy_net.0.weight``, ``action_net.weight`` ...), so the ``policy`` entry of a saved RecurrentPPO / PPO
model loads as it is.  The arithmetic is ``acb200_policy_forward`` (csrc/policy.cu); there is no torch fallback.
"""
from __future__ import annotations

import ctypes as C
from typing import Mapping, Optional

import numpy as np
import torch

from . import _cabi

OBS_DIM, ACT_DIM, HIDDEN, TILE = 18, 4, 256, 128

_LSTM_KEYS = {
    "embed_w": "features_extractor.embedding.0.weight", "embed_b": "features_extractor.embedding.0.bias",
    "l0_w_ih": "features_extractor.lstm.weight_ih_l0", "l0_b_ih": "features_extractor.lstm.bias_ih_l0",
    "l0_b_hh": "features_extractor.lstm.bias_hh_l0",
    "l1_w_ih": "features_extractor.lstm.weight_ih_l1", "l1_b_ih": "features_extractor.lstm.bias_ih_l1",
    "l1_b_hh": "features_extractor.lstm.bias_hh_l1",
    "proj_w": "features_extractor.output_proj.0.weight", "proj_b": "features_extractor.output_proj.0.bias",
    "actor_w_ih": "lstm_actor.weight_ih_l0", "actor_w_hh": "lstm_actor.weight_hh_l0",
    "actor_b_ih": "lstm_actor.bias_ih_l0", "actor_b_hh": "lstm_actor.bias_hh_l0",
    "pi0_w": "mlp_extractor.policy_net.0.weight", "pi0_b": "mlp_extractor.policy_net.0.bias",
    "pi1_w": "mlp_extractor.policy_net.2.weight", "pi1_b": "mlp_extractor.policy_net.2.bias",
    "action_w": "action_net.weight", "action_b": "action_net.bias",
}
_LSTM_SHAPES = {
    "embed_w": (128, 18), "embed_b": (128,), "l0_w_ih": (1024, 128), "l0_b_ih": (1024,), "l0_b_hh": (1024,),
    "l1_w_ih": (1024, 256), "l1_b_ih": (1024,), "l1_b_hh": (1024,), "proj_w": (128, 256), "proj_b": (128,),
    "actor_w_ih": (1024, 128), "actor_w_hh": (1024, 256), "actor_b_ih": (1024,), "actor_b_hh": (1024,),
    "pi0_w": (128, 256), "pi0_b": (128,), "pi1_w": (64, 128), "pi1_b": (64,), "action_w": (4, 64), "action_b": (4,),
}
_MLP_KEYS = {
    "pi0_w": "mlp_extractor.policy_net.0.weight", "pi0_b": "mlp_extractor.policy_net.0.bias",
    "pi1_w": "mlp_extractor.policy_net.2.weight", "pi1_b": "mlp_extractor.policy_net.2.bias",
    "pi2_w": "mlp_extractor.policy_net.4.weight", "pi2_b": "mlp_extractor.policy_net.4.bias",
    "action_w": "action_net.weight", "action_b": "action_net.bias",
}
_MLP_SHAPES = {"pi0_w": (256, 18), "pi0_b": (256,), "pi1_w": (128, 256), "pi1_b": (128,), "pi2_w": (64, 128),
               "pi2_b": (64,), "action_w": (4, 64), "action_b": (4,)}


def _lookup(sd: Mapping[str, torch.Tensor], key: str) -> torch.Tensor:
    # sb3 saves the shared extractor under features_extractor.* and, when not shared, pi_features_extractor.*
    for k in (key, key.replace("features_extractor.", "pi_features_extractor.", 1)):
        if k in sd:
            return sd[k]
    raise KeyError(f"state_dict has no '{key}'")


class LearnedPolicyB200:
    """``kind``: "lstm" (LSTMPolicy / RecurrentPPO) or "mlp" (SimpleMLPPolicy / PPO)."""

    def __init__(self, state_dict: Mapping[str, torch.Tensor], kind: str = "lstm", n: int = 1, device=0,
                 ctx: Optional[_cabi.Context] = None):
        if kind not in ("lstm", "mlp"):
            raise ValueError(f"unknown policy kind {kind!r}")
        self.kind, self.n = kind, int(n)
        self.device = torch.device("cuda", device if isinstance(device, int) else torch.device(device).index or 0)
        self.ctx = ctx or _cabi.Context(self.device.index)
        self.lib = self.ctx.lib
        self.ctx.set_stream(torch.cuda.current_stream(self.device).cuda_stream)
        keys, shapes = (_LSTM_KEYS, _LSTM_SHAPES) if kind == "lstm" else (_MLP_KEYS, _MLP_SHAPES)
        struct = _cabi.LstmPolicyWeightsC() if kind == "lstm" else _cabi.MlpPolicyWeightsC()
        keep = []
        for field, key in keys.items():
            a = np.ascontiguousarray(_lookup(state_dict, key).detach().cpu().float().numpy())
            if a.shape != shapes[field]:
                raise ValueError(f"{key}: expected shape {shapes[field]}, got {a.shape} "
                                 "(architecture other than lstm_policy.py's defaults)")
            keep.append(a)
            setattr(struct, field, a.ctypes.data_as(_cabi._FP))
        h = C.c_void_p()
        create = self.lib.acb200_policy_create_lstm if kind == "lstm" else self.lib.acb200_policy_create_mlp
        self.ctx.check(create(self.ctx.handle, C.byref(struct), C.byref(h)))
        self.handle = h
        hb, cb = C.c_size_t(), C.c_size_t()
        self.ctx.check(self.lib.acb200_policy_state_bytes(self.handle, self.n, C.byref(hb), C.byref(cb)))
        self.h_state = torch.zeros(hb.value, dtype=torch.uint8, device=self.device) if hb.value else None
        self.c_state = torch.zeros(cb.value // 4, dtype=torch.float32, device=self.device) if cb.value else None
        self.action = torch.empty((self.n, ACT_DIM), dtype=torch.float32, device=self.device)

    # ------------------------------------------------------------------------------------------------------------
    @property
    def recurrent(self) -> bool:
        return self.h_state is not None

    def info(self) -> dict:
        f, w, s = C.c_double(), C.c_size_t(), C.c_int()
        self.ctx.check(self.lib.acb200_policy_info(self.handle, C.byref(f), C.byref(w), C.byref(s)))
        return {"flops_per_aircraft": f.value, "weight_bytes": w.value, "smem_bytes": s.value}

    def reset_state(self) -> None:
        if self.recurrent:
            self.h_state.zero_(); self.c_state.zero_()

    def _bind_stream(self):
        """Enqueue on torch's CURRENT stream (as envs._bind_stream does): with a policy-owned context a caller that
        switches streams would otherwise race the producer of obs and the consumer of the action."""
        self.ctx.set_stream(torch.cuda.current_stream(self.device).cuda_stream)

    def _flags(self, t):
        if t is None:
            return None
        t = t if t.dtype == torch.uint8 else t.to(torch.uint8)
        if not (t.is_cuda and t.is_contiguous() and t.numel() == self.n):
            raise ValueError("episode_start must be a contiguous CUDA tensor of n flags")
        return t

    def forward(self, obs: torch.Tensor, episode_start=None, out: Optional[torch.Tensor] = None):
        """obs: CUDA f32 [n, 18] (``BatchedRateControlEnv.obs``); episode_start: CUDA u8/bool [n], a pair of them
        (the env's ``terminated`` and ``truncated``: a fresh episode where either is set) or None.
        Returns the clipped mean action, CUDA f32 [n, 4]; the recurrent state advances in place."""
        if not (obs.is_cuda and obs.dtype == torch.float32 and obs.is_contiguous() and tuple(obs.shape) == (self.n, OBS_DIM)):
            raise ValueError(f"obs must be a contiguous CUDA float32 tensor of shape ({self.n}, {OBS_DIM})")
        if isinstance(episode_start, (tuple, list)):
            es, es_b = self._flags(episode_start[0]), self._flags(episode_start[1])
        else:
            es, es_b = self._flags(episode_start), None
        out = self.action if out is None else out
        if not (out.is_cuda and out.dtype == torch.float32 and out.is_contiguous() and tuple(out.shape) == (self.n, ACT_DIM)):
            raise ValueError(f"out must be a contiguous CUDA float32 tensor of shape ({self.n}, {ACT_DIM})")
        self._bind_stream()
        self.ctx.check(self.lib.acb200_policy_forward(
            self.ctx.handle, self.handle, self.n, obs.data_ptr(), es.data_ptr() if es is not None else None,
            es_b.data_ptr() if es_b is not None else None, self.h_state.data_ptr() if self.recurrent else None, self.c_state.data_ptr() if self.recurrent else None,
            out.data_ptr()))
        return out

    __call__ = forward

    def forward_range(self, obs: torch.Tensor, episode_start, first: int, count: int, out: Optional[torch.Tensor] = None):
        """``forward`` for aircraft [first, first + count) only (``first`` a multiple of 256: whole tile pairs of the
        recurrent-state layout); obs / flags / out are the WHOLE-fleet tensors.  Lets a caller pipeline host copies of one
        part of the fleet behind the forward of another (``step_host_pipelined``).  Runs on torch's current stream."""
        if first % 256 or first < 0 or count < 0 or first + count > self.n:
            raise ValueError("first must be a multiple of 256 and the range must lie inside the fleet")
        if isinstance(episode_start, (tuple, list)):
            es, es_b = self._flags(episode_start[0]), self._flags(episode_start[1])
        else:
            es, es_b = self._flags(episode_start), None
        out = self.action if out is None else out
        self.ctx.set_stream(torch.cuda.current_stream(self.device).cuda_stream)
        tile = first // 128
        self.ctx.check(self.lib.acb200_policy_forward(
            self.ctx.handle, self.handle, count, obs.data_ptr() + first * OBS_DIM * 4,
            es.data_ptr() + first if es is not None else None, es_b.data_ptr() + first if es_b is not None else None,
            self.h_state.data_ptr() + tile * HIDDEN * 128 * 2 if self.recurrent else None,
            self.c_state.data_ptr() + tile * HIDDEN * 128 * 4 if self.recurrent else None,
            out.data_ptr() + first * ACT_DIM * 4))
        return out

    def rollout(self, env, steps: int, parts: int = 0, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """``steps`` closed-loop steps of ``env`` (a ``BatchedRateControlEnv`` with controller "external" or "residual" on the
        SAME context) flown by this policy, in one C-ABI call (``acb200_policy_env_rollout``): per step the action is
        ``forward(env.obs, (env.terminated, env.truncated))`` and the env takes it.  The fleet advances in ``parts``
        independent ranges (0 = automatic) so that one range's env step overlaps another's forward; the results equal the
        step-by-step sequence.  Asynchronous on torch's current stream; capture it in a ``torch.cuda.CUDAGraph`` after one
        warm-up call to replay a whole rollout with one launch.  Returns the action tensor (the last step's actions)."""
        if env.ctx is not self.ctx:
            raise ValueError("policy and env must share one context (pass ctx=env.ctx to LearnedPolicyB200)")
        if env.n != self.n:
            raise ValueError(f"policy is for {self.n} aircraft, env has {env.n}")
        out = self.action if out is None else out
        self._bind_stream()
        self.ctx.check(self.lib.acb200_policy_env_rollout(
            self.ctx.handle, self.handle, C.byref(env.c), C.byref(env.cfg), self.h_state.data_ptr() if self.recurrent else None,
            self.c_state.data_ptr() if self.recurrent else None, out.data_ptr(), int(steps), int(parts)))
        return out

    def predict(self, observation, state=None, episode_start=None, deterministic: bool = True):
        """stable-baselines3's ``predict`` signature over numpy arrays (learned_rate_agent.py:220-227): returns
        ``(actions, state)`` where ``state`` is ``(h, c)`` with shapes (1, n, 256), as sb3-contrib hands them around."""
        if not deterministic:
            raise NotImplementedError("only the deterministic (mean) action is computed on the device")
        obs = torch.as_tensor(np.asarray(observation, np.float32).reshape(self.n, OBS_DIM)).to(self.device)
        if self.recurrent and state is not None:
            self.set_lstm_states(*state)
        es = None if episode_start is None else torch.as_tensor(np.asarray(episode_start).reshape(self.n).astype(np.uint8)).to(self.device)
        act = self.forward(obs, es).cpu().numpy()
        return act, (self.get_lstm_states() if self.recurrent else None)

    # ---- recurrent state as SB3 shapes ---------------------------------------------------------------------------
    def get_lstm_states(self):
        h = torch.empty((self.n, HIDDEN), dtype=torch.float32, device=self.device)
        c = torch.empty_like(h)
        self._bind_stream()
        self.ctx.check(self.lib.acb200_policy_state_export(self.ctx.handle, self.handle, self.n, self.h_state.data_ptr(),
                                                           self.c_state.data_ptr(), h.data_ptr(), c.data_ptr()))
        return h.cpu().numpy()[None], c.cpu().numpy()[None]

    def set_lstm_states(self, h, c) -> None:
        ht = torch.as_tensor(np.asarray(h, np.float32).reshape(self.n, HIDDEN)).to(self.device).contiguous()
        ct = torch.as_tensor(np.asarray(c, np.float32).reshape(self.n, HIDDEN)).to(self.device).contiguous()
        self._bind_stream()
        self.ctx.check(self.lib.acb200_policy_state_import(self.ctx.handle, self.handle, self.n, ht.data_ptr(), ct.data_ptr(),
                                                           self.h_state.data_ptr(), self.c_state.data_ptr()))
        self.ctx.synchronize()

    def close(self) -> None:
        if getattr(self, "handle", None):
            self.lib.acb200_policy_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def random_policy_state_dict(kind: str = "lstm", seed: int = 0) -> dict:
    """Random-init weights of the reference architectures (torch's default initialisers under ``manual_seed(seed)``)
    under stable-baselines3 parameter names -- BASELINE.json C4 ("random-init policy, torch seed 0"); there is no
    checkpoint to load on a machine without network.  Only the initial values are drawn here; no arithmetic of the
    forward path runs in torch."""
    import torch.nn as nn
    g = torch.random.get_rng_state()
    torch.manual_seed(seed)
    sd = {}

    def put(prefix, mod):
        for k, v in mod.state_dict().items():
            sd[f"{prefix}.{k}"] = v.detach().clone()

    if kind == "lstm":
        put("features_extractor.embedding.0", nn.Linear(OBS_DIM, 128))
        put("features_extractor.lstm", nn.LSTM(128, HIDDEN, num_layers=2, batch_first=True))
        put("features_extractor.output_proj.0", nn.Linear(HIDDEN, 128))
        put("lstm_actor", nn.LSTM(128, HIDDEN, num_layers=1))
        put("mlp_extractor.policy_net.0", nn.Linear(HIDDEN, 128))
        put("mlp_extractor.policy_net.2", nn.Linear(128, 64))
        put("action_net", nn.Linear(64, ACT_DIM))
    elif kind == "mlp":
        put("mlp_extractor.policy_net.0", nn.Linear(OBS_DIM, 256))
        put("mlp_extractor.policy_net.2", nn.Linear(256, 128))
        put("mlp_extractor.policy_net.4", nn.Linear(128, 64))
        put("action_net", nn.Linear(64, ACT_DIM))
    else:
        raise ValueError(f"unknown policy kind {kind!r}")
    torch.random.set_rng_state(g)
    return sd


def step_host_pipelined(env, pol: "LearnedPolicyB200", obs_host, reward_host, terminated_host, truncated_host, obs_dev,
                        chunks: int = 4, streams=None):
    """One closed-loop step (policy forward + env step) with the observation coming from and the results going back to
    (pinned) HOST tensors, pipelined over ``chunks`` parts of the fleet on three streams: while part k runs the policy and the
    env kernels, part k+1's observations are on their way in and part k-1's results on their way out.  Returns when the
    host tensors are complete.  Same results as ``obs_dev.copy_(obs_host); env.step(pol(obs_dev, flags))`` + copies out."""
    n = env.n
    per = ((n + chunks - 1) // chunks + 255) // 256 * 256
    dev = env.device
    if streams is None:
        streams = getattr(env, "_pipe_streams", None)
        if streams is None:
            streams = env._pipe_streams = tuple(torch.cuda.Stream(dev) for _ in range(3))
    s_in, s_run, s_out = streams
    flags = (env.terminated, env.truncated)
    start = torch.cuda.current_stream(dev).record_event()
    for s_ in streams:
        s_.wait_event(start)
    reward = env.reward
    for lo in range(0, n, per):
        cnt = min(per, n - lo)
        with torch.cuda.stream(s_in):
            obs_dev[lo:lo + cnt].copy_(obs_host[lo:lo + cnt], non_blocking=True)
            ready = s_in.record_event()
        with torch.cuda.stream(s_run):
            s_run.wait_event(ready)
            pol.forward_range(obs_dev, flags, lo, cnt)
            env._bind_stream()
            env.ctx.check(env.ctx.lib.acb200_env_step_range(env.ctx.handle, C.byref(env.c), C.byref(env.cfg),
                                                            C.c_void_p(pol.action.data_ptr()), lo, cnt))
            done = s_run.record_event()
        with torch.cuda.stream(s_out):
            s_out.wait_event(done)
            obs_host[lo:lo + cnt].copy_(env.obs[lo:lo + cnt], non_blocking=True)
            reward_host[lo:lo + cnt].copy_(reward[lo:lo + cnt], non_blocking=True)
            terminated_host[lo:lo + cnt].copy_(env.terminated[lo:lo + cnt], non_blocking=True)
            truncated_host[lo:lo + cnt].copy_(env.truncated[lo:lo + cnt], non_blocking=True)
    s_out.synchronize()
    torch.cuda.current_stream(dev).wait_stream(s_run)
    return obs_host, reward_host, terminated_host, truncated_host
