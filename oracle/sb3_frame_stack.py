"""CPU restatement of stable_baselines3's VecFrameStack for 1-D observations.

TEST INFRASTRUCTURE ONLY (tests/ may import it; the product never does).

stable_baselines3 is a third-party dependency of the reference's training scripts (train_test/sb3_train_vec.py:3-8,70:
``VecFrameStack(env, n_stack=5)``) and is neither vendored in /root/reference nor installed here; no version is pinned
anywhere in the reference (setup.py:5 lists only gym).  This restates the published algorithm of
``stable_baselines3.common.vec_env.stacked_observations.StackedObservations`` (v1.8 - v2.x, unchanged for Box
observations): ``reset`` zero-fills and stores the observation in the last slot; ``update`` rolls by one frame, and
for a done env builds the stacked terminal observation from the rolled stack's first n_stack - 1 frames plus the
env's terminal observation, then zero-fills that env's stack before the new (reset) observation is stored.
Parity is unpinned against SB3 itself (absent); it is pinned to this restatement.
"""
import numpy as np


class StackedObservations:
    def __init__(self, num_envs, n_stack, obs_dim):
        self.n_stack, self.obs_dim = n_stack, obs_dim
        self.stacked_obs = np.zeros((num_envs, n_stack * obs_dim), dtype=np.float32)

    def reset(self, observation):
        self.stacked_obs[...] = 0
        self.stacked_obs[..., -self.obs_dim:] = observation
        return self.stacked_obs

    def update(self, observations, dones, infos):
        shift = -self.obs_dim
        self.stacked_obs = np.roll(self.stacked_obs, shift, axis=-1)
        for env_idx, done in enumerate(dones):
            if done:
                if "terminal_observation" in infos[env_idx]:
                    old_terminal = infos[env_idx]["terminal_observation"]
                    previous_stack = self.stacked_obs[env_idx, ..., :shift]
                    infos[env_idx]["terminal_observation"] = np.concatenate((previous_stack, old_terminal), axis=-1)
                self.stacked_obs[env_idx] = 0
        self.stacked_obs[..., shift:] = observations
        return self.stacked_obs, infos
