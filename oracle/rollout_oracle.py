"""TEST INFRASTRUCTURE -- numpy restatement of stable-baselines3's ``RolloutBuffer.compute_returns_and_advantage``
(third party, ``stable-baselines3>=2.0.0`` in the reference's requirements.txt:26, not installed here and not under
/root/reference), the step the reference's PPO trainers run after every rollout
(learned_controllers/utils/training_utils.py:112-160, train_overnight.py:151-167).

Restated from the published algorithm (GAE, Schulman et al. 2016, with SB3's episode_start convention), float32 arrays
and Python-float coefficients exactly as SB3 holds them, so numpy's own casting rules decide the rounding.
PARITY UNPINNED against SB3 itself (no fixture can be generated offline); pinned against a scalar float64 evaluation
of the recursion in tests/test_rollout.py.  Only tests/ and bench baselines may import this module.
"""
import numpy as np


def compute_returns_and_advantage(rewards, values, episode_starts, last_values, dones, gamma=0.99, gae_lambda=0.95):
    """rewards, values, episode_starts: float32 [T][n]; last_values float32 [n]; dones bool [n] -> (advantages, returns)."""
    rewards = np.asarray(rewards, np.float32); values = np.asarray(values, np.float32)
    episode_starts = np.asarray(episode_starts, np.float32)
    last_values = np.asarray(last_values, np.float32).copy().flatten()
    T = rewards.shape[0]
    advantages = np.zeros_like(rewards)
    last_gae_lam = 0
    for step in reversed(range(T)):
        if step == T - 1:
            next_non_terminal = 1.0 - np.asarray(dones).astype(np.float32)
            next_values = last_values
        else:
            next_non_terminal = 1.0 - episode_starts[step + 1]
            next_values = values[step + 1]
        delta = rewards[step] + gamma * next_values * next_non_terminal - values[step]
        last_gae_lam = delta + gamma * gae_lambda * next_non_terminal * last_gae_lam
        advantages[step] = last_gae_lam
    returns = advantages + values
    return advantages, returns
