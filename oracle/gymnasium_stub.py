"""Writes a minimal `gymnasium` stand-in into oracle/_ref/gymnasium (TEST INFRASTRUCTURE ONLY).

gymnasium is not installed in this image and cannot be (no network).  The reference's
learned_controllers/envs/rate_env.py only needs `gym.Env`, `gym.spaces.Box` and `gym.register`
to import and run unmodified; this provides exactly those.
"""
import os
import sys

INIT = '''"""gymnasium stand-in for running the reference env as an oracle (see oracle/gymnasium_stub.py)."""
from . import spaces  # noqa: F401


class Env:
    metadata = {}

    def reset(self, seed=None, options=None):
        return None


def register(*args, **kwargs):
    return None
'''

SPACES = '''import numpy as np


class Box:
    def __init__(self, low, high, shape=None, dtype=np.float32):
        self.dtype = np.dtype(dtype)
        self.low = np.asarray(low, dtype=self.dtype)
        self.high = np.asarray(high, dtype=self.dtype)
        self.shape = self.low.shape if shape is None else tuple(shape)
'''

if __name__ == "__main__":
    out = sys.argv[1]
    os.makedirs(out, exist_ok=True)
    with open(os.path.join(out, "__init__.py"), "w") as f:
        f.write(INIT)
    with open(os.path.join(out, "spaces.py"), "w") as f:
        f.write(SPACES)
