"""CPU restatement of stable_baselines3 ``ReplayBuffer`` (add / sample) and ``OffPolicyAlgorithm._store_transition``.

TEST INFRASTRUCTURE ONLY.  stable_baselines3 is an un-vendored, un-pinned third-party dependency of the reference's
training scripts (train_test/sb3_train_vec.py:3-8, :72 SAC); this restates its published algorithm
(common/buffers.py ``ReplayBuffer`` with optimize_memory_usage=False and handle_timeout_termination=True,
common/off_policy_algorithm.py ``_store_transition``) in numpy.  Parity unpinned against SB3 itself (absent here).
"""
import numpy as np


class ReplayBuffer:
    def __init__(self, buffer_size, n_envs, obs_dim, act_dim):
        self.buffer_size = max(buffer_size // n_envs, 1)
        self.n_envs = n_envs
        self.observations = np.zeros((self.buffer_size, n_envs, obs_dim), np.float32)
        self.next_observations = np.zeros((self.buffer_size, n_envs, obs_dim), np.float32)
        self.actions = np.zeros((self.buffer_size, n_envs, act_dim), np.float32)
        self.rewards = np.zeros((self.buffer_size, n_envs), np.float32)
        self.dones = np.zeros((self.buffer_size, n_envs), np.float32)
        self.timeouts = np.zeros((self.buffer_size, n_envs), np.float32)
        self.pos, self.full = 0, False

    def add(self, obs, next_obs, action, reward, done, infos):
        self.observations[self.pos] = np.array(obs)
        self.next_observations[self.pos] = np.array(next_obs)
        self.actions[self.pos] = np.array(action).reshape(self.n_envs, -1)
        self.rewards[self.pos] = np.array(reward)
        self.dones[self.pos] = np.array(done)
        self.timeouts[self.pos] = np.array([info.get("TimeLimit.truncated", False) for info in infos])
        self.pos += 1
        if self.pos == self.buffer_size:
            self.full, self.pos = True, 0

    def get_samples(self, batch_inds, env_indices):
        return dict(
            observations=self.observations[batch_inds, env_indices, :],
            actions=self.actions[batch_inds, env_indices, :],
            next_observations=self.next_observations[batch_inds, env_indices, :],
            dones=(self.dones[batch_inds, env_indices] * (1 - self.timeouts[batch_inds, env_indices])).reshape(-1, 1),
            rewards=self.rewards[batch_inds, env_indices].reshape(-1, 1))


def store_transition(buf, last_obs, new_obs, actions, rewards, dones, infos):
    """_store_transition: the next observation of a finished env is infos[i]["terminal_observation"]."""
    next_obs = np.array(new_obs, copy=True)
    for i, done in enumerate(dones):
        if done and infos[i].get("terminal_observation") is not None:
            next_obs[i] = infos[i]["terminal_observation"]
    buf.add(last_obs, next_obs, actions, rewards, dones, infos)
