"""Writes a minimal `stable_baselines3` stand-in into oracle/_ref/stable_baselines3 (TEST INFRASTRUCTURE ONLY).

stable-baselines3 is not installed in this image and cannot be (no network).  The reference's
learned_controllers/networks/lstm_policy.py needs one name from it to import unmodified (and tensorboard_flight/callbacks.py a second, common.callbacks.BaseCallback):
`stable_baselines3.common.torch_layers.BaseFeaturesExtractor` (lstm_policy.py:9,13), whose published behaviour is an
`nn.Module` that stores the observation space and `features_dim` (its `__init__(observation_space, features_dim)` asserts
features_dim > 0 and exposes the `features_dim` property).  This provides that class and nothing else, so that
tests/golden/make_golden.py can run the reference's own LSTMFeaturesExtractor and pin oracle/policy_oracle.py on it.
"""
import os
import sys

INIT = '"""stable_baselines3 stand-in: only common.torch_layers.BaseFeaturesExtractor (see oracle/sb3_stub.py)."""\n'
COMMON = ""
TORCH_LAYERS = '''import torch.nn as nn


class BaseFeaturesExtractor(nn.Module):
    """stable_baselines3.common.torch_layers.BaseFeaturesExtractor, as published: stores the space and features_dim."""

    def __init__(self, observation_space, features_dim: int = 0):
        super().__init__()
        assert features_dim > 0
        self._observation_space = observation_space
        self._features_dim = features_dim

    @property
    def features_dim(self) -> int:
        return self._features_dim
'''

CALLBACKS = '''class BaseCallback:
    """stable_baselines3.common.callbacks.BaseCallback, the part FlightLoggerCallback uses: verbose, locals, num_timesteps."""

    def __init__(self, verbose: int = 0):
        self.verbose = verbose
        self.locals = {}
        self.globals = {}
        self.num_timesteps = 0
        self.n_calls = 0
        self.model = None

    def on_step(self) -> bool:
        self.n_calls += 1
        return self._on_step()
'''

if __name__ == "__main__":
    out = sys.argv[1]
    os.makedirs(os.path.join(out, "common"), exist_ok=True)
    open(os.path.join(out, "__init__.py"), "w").write(INIT)
    open(os.path.join(out, "common", "__init__.py"), "w").write(COMMON)
    open(os.path.join(out, "common", "torch_layers.py"), "w").write(TORCH_LAYERS)
    open(os.path.join(out, "common", "callbacks.py"), "w").write(CALLBACKS)
    open(os.path.join(out, "common", "vec_env.py"), "w").write('class VecEnv:\n    """name only: tensorboard_flight/callbacks.py imports it for a type annotation"""\n')
