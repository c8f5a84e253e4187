"""CPU restatement of stable_baselines3 ``RolloutBuffer.compute_returns_and_advantage`` (GAE(lambda)).

TEST INFRASTRUCTURE ONLY.  stable_baselines3 is an un-vendored, un-pinned third-party dependency of the reference's
training scripts (train_test/sb3_train_vec.py:3-8); this restates its published algorithm (common/buffers.py,
unchanged v1.0 - v2.x) in float32 numpy, operation for operation: gamma and gae_lambda stay Python floats as in
SB3, so numpy rounds gamma to float32 against the arrays and takes gamma * gae_lambda in float64 before it meets a
float32 array.  Parity unpinned against SB3 itself (absent here).
"""
import numpy as np


def compute_returns_and_advantage(rewards, values, episode_starts, last_values, dones, gamma=0.99, gae_lambda=0.95):
    rewards, values, episode_starts = (np.asarray(a, dtype=np.float32) for a in (rewards, values, episode_starts))
    last_values = np.asarray(last_values, dtype=np.float32)
    gamma, gae_lambda = float(gamma), float(gae_lambda)
    buffer_size = rewards.shape[0]
    advantages = np.zeros_like(rewards)
    last_gae_lam = np.float32(0)
    for step in reversed(range(buffer_size)):
        if step == buffer_size - 1:
            next_non_terminal = np.float32(1.0) - np.asarray(dones).astype(np.float32)
            next_values = last_values
        else:
            next_non_terminal = np.float32(1.0) - episode_starts[step + 1]
            next_values = values[step + 1]
        delta = rewards[step] + gamma * next_values * next_non_terminal - values[step]
        last_gae_lam = delta + gamma * gae_lambda * next_non_terminal * last_gae_lam
        advantages[step] = last_gae_lam
    return advantages, advantages + values
