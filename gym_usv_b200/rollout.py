"""Tensor-native on-policy rollout collection against the GPU VecEnv (BASELINE config 5).

Stock stable_baselines3 ``OnPolicyAlgorithm.collect_rollouts`` moves every observation to the device
(``obs_as_tensor``) and every action back (``.cpu().numpy()``) once per step; with the env itself on the GPU those
round trips are the whole cost.  ``collect_rollouts`` below is the same loop -- policy forward, action clip, env step,
time-limit bootstrap of truncated episodes, buffer add, then GAE -- with every tensor staying on the device:
observations, actions, rewards, episode starts, values and log-probabilities are written straight into a device
``[n_steps, N, ...]`` buffer, and ``RolloutBuffer.compute_returns_and_advantage`` is one kernel (``usvgpu_gae``).

The shapes follow the reference's PPO settings (train_test/config.py:3-15: n_steps 2048, net_arch pi/vf [256, 256],
ReLU, log_std_init -2).  The policy is plain torch (library GEMMs): it is the caller's model, not part of the env path.
SB3 itself is absent from this image; where it is installed, the same function can be bound as
``PPO.collect_rollouts``.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import torch
from torch import nn

from . import _native as nat


class MlpActorCritic(nn.Module):
    """SB3 ``ActorCriticPolicy`` shape for Box actions: separate pi / vf towers, state-independent log_std."""

    def __init__(self, obs_dim: int, act_dim: int, hidden=(256, 256), log_std_init: float = -2.0):
        super().__init__()

        def tower(out):
            layers, d = [], obs_dim
            for h in hidden:
                layers += [nn.Linear(d, h), nn.ReLU()]
                d = h
            return nn.Sequential(*layers, nn.Linear(d, out))

        self.pi, self.vf = tower(act_dim), tower(1)
        self.log_std = nn.Parameter(torch.full((act_dim,), float(log_std_init)))

    @torch.no_grad()
    def forward(self, obs: torch.Tensor):
        """-> (actions, values, log_probs), sampled from the diagonal Gaussian."""
        mean = self.pi(obs)
        std = self.log_std.exp()
        noise = torch.randn_like(mean)
        actions = mean + std * noise
        log_prob = (-0.5 * noise.pow(2) - self.log_std - 0.9189385332046727).sum(-1)
        return actions, self.vf(obs).squeeze(-1), log_prob

    @torch.no_grad()
    def predict_values(self, obs: torch.Tensor) -> torch.Tensor:
        return self.vf(obs).squeeze(-1)


@dataclass
class DeviceRolloutBuffer:
    """stable_baselines3 ``RolloutBuffer`` fields as device tensors ``[n_steps, N, ...]``."""
    observations: torch.Tensor
    actions: torch.Tensor
    rewards: torch.Tensor
    episode_starts: torch.Tensor
    values: torch.Tensor
    log_probs: torch.Tensor
    advantages: torch.Tensor
    returns: torch.Tensor

    @classmethod
    def allocate(cls, n_steps: int, n_envs: int, obs_dim: int, act_dim: int, device) -> "DeviceRolloutBuffer":
        z = lambda *shape: torch.zeros(shape, dtype=torch.float32, device=device)  # noqa: E731
        return cls(z(n_steps, n_envs, obs_dim), z(n_steps, n_envs, act_dim), z(n_steps, n_envs), z(n_steps, n_envs),
                   z(n_steps, n_envs), z(n_steps, n_envs), z(n_steps, n_envs), z(n_steps, n_envs))

    def compute_returns_and_advantage(self, last_values: torch.Tensor, dones: torch.Tensor, gamma: float = 0.99,
                                      gae_lambda: float = 0.95) -> None:
        """RolloutBuffer.compute_returns_and_advantage as one kernel (usvgpu_gae)."""
        n_steps, n = self.rewards.shape
        dev = self.rewards.device
        last_values = last_values.to(torch.float32).contiguous()
        dones = dones.to(torch.uint8).contiguous()
        p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
        nat.check(nat.lib().usvgpu_gae(nat.device_index(dev), n_steps, n, p(self.rewards), p(self.values),
                                       p(self.episode_starts), p(last_values), p(dones), float(gamma),
                                       float(gae_lambda), p(self.advantages), p(self.returns),
                                       C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)))


def collect_rollouts(env, policy, buf: DeviceRolloutBuffer, last_obs: torch.Tensor, last_episode_starts: torch.Tensor,
                     gamma: float = 0.99, gae_lambda: float = 0.95, step_policy: bool = True):
    """OnPolicyAlgorithm.collect_rollouts with device-resident tensors.

    env: UsvGpuVecEnv or UsvGpuVecFrameStack (``step_tensor`` face).  Returns (last_obs, last_episode_starts) for the
    next call.  step_policy=False skips the policy forward (zero-mean unit actions are replayed from the buffer) so
    that the env's share of the rollout time can be measured.
    """
    n_steps = buf.rewards.shape[0]
    venv = env
    while hasattr(venv, "venv"):  # UsvGpuVecFrameStack / UsvGpuVecVideoRecorder wrappers
        venv = venv.venv
    low = torch.as_tensor(venv.action_space.low, device=last_obs.device)
    high = torch.as_tensor(venv.action_space.high, device=last_obs.device)
    stacked = hasattr(env, "terminal_stack")
    values = None
    for t in range(n_steps):
        if step_policy:
            actions, values, log_probs = policy(last_obs)
        else:
            actions, values, log_probs = buf.actions[t], buf.values[t], buf.log_probs[t]
        buf.observations[t].copy_(last_obs)
        buf.actions[t] = actions
        buf.values[t] = values
        buf.log_probs[t] = log_probs
        buf.episode_starts[t] = last_episode_starts
        clipped = torch.clamp(actions, low, high).contiguous()  # SB3 clips Box actions before env.step
        new_obs, rewards, term, trunc = env.step_tensor(clipped)
        # time-limit bootstrap (SB3: rewards[idx] += gamma * V(terminal_observation) for TimeLimit.truncated)
        tl = trunc.bool() & ~term.bool()
        if step_policy:
            tobs = env.terminal_stack if stacked else venv.terminal_obs
            tv = policy.predict_values(tobs)  # whole batch: cheaper than a data-dependent gather + sync
            rewards = torch.where(tl, rewards + gamma * tv, rewards)
        buf.rewards[t] = rewards
        last_obs = new_obs
        last_episode_starts = (term | trunc).to(torch.float32)
    last_values = policy.predict_values(last_obs) if step_policy else values
    buf.compute_returns_and_advantage(last_values, last_episode_starts, gamma, gae_lambda)
    return last_obs, last_episode_starts
