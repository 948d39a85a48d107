"""Shared helpers for the parity tests."""
import numpy as np


def sample_initial_states(rng, n):
    """SURVEY.md 8(d) C2 inputs: V~U(15,30), h~U(50,200), phi,theta~U(+-15deg), psi~U(0,2pi), pqr~U(+-0.1)."""
    s = np.zeros((13, n))
    s[2] = -rng.uniform(50, 200, n)
    s[3] = rng.uniform(15, 30, n)
    s[6] = rng.uniform(-np.radians(15), np.radians(15), n)
    s[7] = rng.uniform(-np.radians(15), np.radians(15), n)
    s[8] = rng.uniform(0, 2 * np.pi, n)
    s[9:12] = rng.uniform(-0.1, 0.1, (3, n))
    return s


def surface_sequence(rng, n, nseg, amp_cycle=(0.01, 0.01, 0.03, 0.01, 0.1, 0.01, 0.03, 0.3)):
    """Piecewise-constant surfaces, per-aircraft amplitude cycling from gentle to violent."""
    amp = np.resize(np.array(amp_cycle), n)
    ctrl = np.zeros((nseg, 4, n))
    ctrl[:, 0:3] = rng.uniform(-1.0, 1.0, (nseg, 3, n)) * amp
    ctrl[:, 3] = rng.uniform(0.3, 0.9, (nseg, n))
    return ctrl


def in_envelope(state, derived):
    """not RateControlEnv._check_termination (rate_env.py:437-460); state [13][n], derived [4][n]."""
    return ~((derived[1] < 5.0) | (np.abs(state[6]) > np.radians(120)) | (np.abs(state[7]) > np.radians(80))
             | (derived[0] < 8.0))


def rel_err_cols(a, b):
    """per-aircraft max |a-b| / max(|b|,1) over the fields (axis 0)."""
    return np.max(np.abs(a - b) / np.maximum(np.abs(b), 1.0), axis=0)
