"""TEST INFRASTRUCTURE -- CPU/torch fp32 restatement of the learned rate policy's actor forward.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s baseline legs may import this module; the product
path (``acb200_policy_forward``, csrc/policy.cu) never does.

What it restates
  * ``LSTMFeaturesExtractor`` -- learned_controllers/networks/lstm_policy.py:13-97: Linear(18,128)+ReLU ->
    nn.LSTM(128, 256, num_layers=2, batch_first=True) called WITHOUT state on a length-1 sequence (so h0=c0=0 on every
    call, :86) -> Linear(256,128)+ReLU.  Layer names (``embedding.0``, ``lstm``, ``output_proj.0``) are the
    reference's, so a reference ``state_dict`` loads.
  * ``LSTMPolicy.get_policy_kwargs`` -- lstm_policy.py:107-136: ``net_arch=dict(pi=[128, 64], vf=[128, 64])``, ReLU.
  * ``SimpleMLPPolicy.get_policy_kwargs`` -- lstm_policy.py:145-164: pi=[256, 128, 64], ReLU, flatten extractor.
  * The stateful actor LSTM, the pi-MLP, the Gaussian mean head and the action clipping are third-party
    (sb3-contrib ``RecurrentActorCriticPolicy`` / stable-baselines3 ``ActorCriticPolicy``, requirements.txt:26-27,
    ``>=2.0.0`` unpinned, NOT under /root/reference).  Restated from their published behaviour: ``lstm_actor =
    nn.LSTM(features_dim, 256, num_layers=1)``; states are multiplied by ``(1 - episode_start)`` before the step;
    ``mlp_extractor.policy_net`` = Linear/ReLU stack; ``action_net = Linear(last, 4)``; ``predict(deterministic=True)``
    returns the mean clipped to the action space (call sites: controllers/learned_rate_agent.py:220-227).

PARITY UNPINNED: the reference holds no golden vector for this path and stable-baselines3 / sb3-contrib are not
installed here, so this restatement is checked only against torch's own nn.LSTM / nn.Linear (which it is built
from).  The CUDA kernel computes with bf16 operands and fp32 accumulation; the tolerance is stated in
tests/test_gpu_policy.py.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn

OBS_DIM, ACT_DIM, LSTM_HIDDEN, FEATURES_DIM = 18, 4, 256, 128
ACTION_LOW = torch.tensor([-1.0, -1.0, -1.0, 0.0])      # learned_controllers/envs/rate_env.py:128-138
ACTION_HIGH = torch.tensor([1.0, 1.0, 1.0, 1.0])


class LSTMFeaturesExtractorOracle(nn.Module):
    """lstm_policy.py:13-97 with ``BaseFeaturesExtractor`` replaced by ``nn.Module`` (it only stores features_dim)."""

    def __init__(self, obs_dim: int = OBS_DIM, features_dim: int = FEATURES_DIM, lstm_hidden_size: int = LSTM_HIDDEN,
                 n_lstm_layers: int = 2):
        super().__init__()
        self.features_dim = features_dim
        self.embedding = nn.Sequential(nn.Linear(obs_dim, 128), nn.ReLU())
        self.lstm = nn.LSTM(input_size=128, hidden_size=lstm_hidden_size, num_layers=n_lstm_layers, batch_first=True)
        self.output_proj = nn.Sequential(nn.Linear(lstm_hidden_size, features_dim), nn.ReLU())

    def forward(self, observations: torch.Tensor) -> torch.Tensor:
        squeeze = observations.dim() == 2
        if squeeze:
            observations = observations.unsqueeze(1)
        lstm_out, _ = self.lstm(self.embedding(observations))          # no state passed: zeros every call
        lstm_out = lstm_out.squeeze(1) if squeeze else lstm_out[:, -1, :]
        return self.output_proj(lstm_out)


class _PolicyNet(nn.Module):
    def __init__(self, sizes):
        super().__init__()
        layers = []
        for a, b in zip(sizes[:-1], sizes[1:]):
            layers += [nn.Linear(a, b), nn.ReLU()]
        self.policy_net = nn.Sequential(*layers)                      # names: mlp_extractor.policy_net.{0,2,4}


class RecurrentActorOracle(nn.Module):
    """Actor side of sb3-contrib's recurrent policy as ``LSTMPolicy.get_policy_kwargs()`` configures it."""

    kind = "lstm"

    def __init__(self):
        super().__init__()
        self.features_extractor = LSTMFeaturesExtractorOracle()
        self.lstm_actor = nn.LSTM(FEATURES_DIM, LSTM_HIDDEN, num_layers=1)
        self.mlp_extractor = _PolicyNet([LSTM_HIDDEN, 128, 64])
        self.action_net = nn.Linear(64, ACT_DIM)
        self.log_std = nn.Parameter(torch.zeros(ACT_DIM))                # unused by the deterministic forward

    def initial_state(self, n: int):
        z = torch.zeros(n, LSTM_HIDDEN)
        return z, z.clone()

    @torch.no_grad()
    def forward(self, obs: torch.Tensor, state, episode_start: torch.Tensor | None = None):
        """obs [n,18] f32, state (h [n,256], c [n,256]), episode_start [n] bool -> (action [n,4], (h', c'))."""
        h, c = state
        if episode_start is not None:
            keep = (~episode_start.bool()).to(h.dtype).unsqueeze(1)
            h, c = h * keep, c * keep
        feats = self.features_extractor(obs)
        out, (h2, c2) = self.lstm_actor(feats.unsqueeze(0), (h.unsqueeze(0), c.unsqueeze(0)))
        latent = self.mlp_extractor.policy_net(out.squeeze(0))
        mean = self.action_net(latent)
        return torch.max(torch.min(mean, ACTION_HIGH), ACTION_LOW), (h2.squeeze(0), c2.squeeze(0))


class MlpActorOracle(nn.Module):
    """Actor side of stable-baselines3's ``MlpPolicy`` with ``SimpleMLPPolicy.get_policy_kwargs()``."""

    kind = "mlp"

    def __init__(self, net_arch=(256, 128, 64)):
        super().__init__()
        self.mlp_extractor = _PolicyNet([OBS_DIM, *net_arch])
        self.action_net = nn.Linear(net_arch[-1], ACT_DIM)
        self.log_std = nn.Parameter(torch.zeros(ACT_DIM))

    def initial_state(self, n: int):
        return None

    @torch.no_grad()
    def forward(self, obs: torch.Tensor, state=None, episode_start=None):
        mean = self.action_net(self.mlp_extractor.policy_net(obs))
        return torch.max(torch.min(mean, ACTION_HIGH), ACTION_LOW), None


def make_policy(kind: str = "lstm", seed: int = 0, scale: float = 1.0) -> nn.Module:
    """Random-init policy (torch default init under ``torch.manual_seed(seed)``; BASELINE C4 says seed 0).
    ``scale`` multiplies every weight matrix: >1 drives the gates out of their linear range for harder parity cases."""
    g = torch.random.get_rng_state()
    torch.manual_seed(seed)
    net = RecurrentActorOracle() if kind == "lstm" else MlpActorOracle()
    torch.random.set_rng_state(g)
    if scale != 1.0:
        with torch.no_grad():
            for name, p in net.named_parameters():
                if "weight" in name:
                    p.mul_(scale)
    return net.eval()


def synthetic_obs(n: int, seed: int = 0) -> torch.Tensor:
    """Observations with the magnitudes RateControlEnv produces (rate_env.py:318-372): rates/errors O(1), airspeed
    ~20, altitude ~100, Euler angles, previous action."""
    r = np.random.default_rng(seed)
    o = np.empty((n, OBS_DIM), np.float32)
    o[:, 0:9] = r.normal(0.0, 0.6, (n, 9))
    o[:, 9] = r.uniform(12.0, 30.0, n)
    o[:, 10] = r.uniform(40.0, 160.0, n)
    o[:, 11:14] = r.uniform(-1.0, 1.0, (n, 3))
    o[:, 14:17] = r.uniform(-1.0, 1.0, (n, 3))
    o[:, 17] = r.uniform(0.0, 1.0, n)
    return torch.from_numpy(o)


# ---- the same forward with the kernel's operand rounding (bf16 operands, fp32 accumulate / gates) -------------------
def _bf(x: torch.Tensor) -> torch.Tensor:
    return x.to(torch.bfloat16).to(torch.float32)


def _lin(x, w, b):
    return _bf(x).double() @ _bf(w).double().t() + b.double()


def _lstm_cell(x, h, c, w_ih, w_hh, b_ih, b_hh):
    g = _bf(x).double() @ _bf(w_ih).double().t() + (b_ih + b_hh).double()
    if h is not None:
        g = g + _bf(h).double() @ _bf(w_hh).double().t()
    i, f, gg, o = g.float().chunk(4, dim=1)
    c2 = torch.sigmoid(i) * torch.tanh(gg) + (torch.sigmoid(f) * c if c is not None else 0.0)
    return torch.sigmoid(o) * torch.tanh(c2), c2


@torch.no_grad()
def forward_bf16_emulated(net: nn.Module, obs: torch.Tensor, state=None, episode_start=None):
    """What csrc/policy.cu computes, in torch: every GEMM operand rounded to bf16 (observation and first-layer
    weights as hi + lo pairs, i.e. 16 mantissa bits each), products accumulated exactly (float64 here, fp32 in tensor memory), bias / gates / ReLU in
    fp32, h carried as bf16 and c as fp32.  Separates 'the kernel is wrong' from 'bf16 is coarse' in the tests."""
    relu = torch.relu

    def first(w, b):     # observation and first-layer weight as hi + lo bf16 pairs; the lo*lo product is dropped
        o_hi, w_hi = _bf(obs), _bf(w)
        o_lo, w_lo = _bf(obs - o_hi), _bf(w - w_hi)
        return ((o_hi + o_lo).double() @ w_hi.double().t() + o_hi.double() @ w_lo.double().t() + b.double()).float()

    if net.kind == "mlp":
        p = net.mlp_extractor.policy_net
        x = relu(first(p[0].weight, p[0].bias))
        x = relu(_lin(x, p[2].weight, p[2].bias).float())
        x = relu(_lin(x, p[4].weight, p[4].bias).float())
        mean = _lin(x, net.action_net.weight, net.action_net.bias).float()
        return torch.max(torch.min(mean, ACTION_HIGH), ACTION_LOW), None
    fe = net.features_extractor
    x = relu(first(fe.embedding[0].weight, fe.embedding[0].bias))
    x, _ = _lstm_cell(x, None, None, fe.lstm.weight_ih_l0, None, fe.lstm.bias_ih_l0, fe.lstm.bias_hh_l0)
    x, _ = _lstm_cell(x, None, None, fe.lstm.weight_ih_l1, None, fe.lstm.bias_ih_l1, fe.lstm.bias_hh_l1)
    x = relu(_lin(x, fe.output_proj[0].weight, fe.output_proj[0].bias).float())
    h, c = state
    if episode_start is not None:
        keep = (~episode_start.bool()).float().unsqueeze(1)
        h, c = h * keep, c * keep
    la = net.lstm_actor
    h2, c2 = _lstm_cell(x, h, c, la.weight_ih_l0, la.weight_hh_l0, la.bias_ih_l0, la.bias_hh_l0)
    h2 = _bf(h2)
    p = net.mlp_extractor.policy_net
    x = relu(_lin(h2, p[0].weight, p[0].bias).float())
    x = relu(_lin(x, p[2].weight, p[2].bias).float())
    mean = _lin(x, net.action_net.weight, net.action_net.bias).float()
    return torch.max(torch.min(mean, ACTION_HIGH), ACTION_LOW), (h2, c2)
