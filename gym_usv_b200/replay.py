"""Device-resident replay buffer and transition collection for off-policy learners.

``train_test/sb3_train_vec.py:72`` trains **SAC** with ``config_sac`` (``train_test/config.py:17-37``: buffer_size
400 000, batch_size 256, train_freq 8, gradient_steps 8, learning_starts 50 000).  Stock stable_baselines3 keeps the
replay buffer in host numpy arrays: every env step ends with ``ReplayBuffer.add`` on host copies of the observations,
and every gradient step starts with ``sample`` + ``to_torch``.  With the env on the GPU both directions are pure
PCIe traffic.  ``DeviceReplayBuffer`` keeps SB3's layout (``[capacity, n_envs, ...]``, ``pos`` / ``full``,
``dones`` + ``timeouts``) in HBM: ``add`` is one kernel (``usvgpu_replay_add``: writes the slot of all envs, picks
the terminal observation for finished envs), ``sample`` one gather kernel (``usvgpu_replay_sample``).

SB3 itself is absent from this image; the checker is the numpy restatement ``oracle/sb3_replay.py`` (tests only).
"""
from __future__ import annotations

import ctypes as C
from typing import NamedTuple, Optional

import torch

from . import _native as nat


class ReplaySamples(NamedTuple):  # stable_baselines3.common.type_aliases.ReplayBufferSamples
    observations: torch.Tensor
    actions: torch.Tensor
    next_observations: torch.Tensor
    dones: torch.Tensor
    rewards: torch.Tensor


class DeviceReplayBuffer:
    """stable_baselines3 ``ReplayBuffer`` (optimize_memory_usage=False, handle_timeout_termination=True) in HBM."""

    def __init__(self, buffer_size: int, n_envs: int, obs_dim: int, act_dim: int, device):
        self.device = torch.device(device)
        self.n_envs, self.obs_dim, self.act_dim = int(n_envs), int(obs_dim), int(act_dim)
        self.buffer_size = max(int(buffer_size) // self.n_envs, 1)  # SB3: per-env capacity
        z = lambda shape, dt: torch.zeros(shape, dtype=dt, device=self.device)  # noqa: E731
        cap, n = self.buffer_size, self.n_envs
        self.observations = z((cap, n, obs_dim), torch.float32)
        self.next_observations = z((cap, n, obs_dim), torch.float32)
        self.actions = z((cap, n, act_dim), torch.float32)
        self.rewards = z((cap, n), torch.float32)
        self.dones = z((cap, n), torch.uint8)
        self.timeouts = z((cap, n), torch.uint8)
        self.pos, self.full = 0, False

    def size(self) -> int:
        return self.buffer_size if self.full else self.pos

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    @staticmethod
    def _rows(t: torch.Tensor, dim: int):
        """(pointer, row stride in elements) of a [N, dim] float32 tensor with unit inner stride (views allowed)."""
        assert t.dtype == torch.float32 and t.dim() == 2 and t.shape[1] == dim and (dim == 1 or t.stride(1) == 1)
        return C.c_void_p(t.data_ptr()), int(t.stride(0))

    def add(self, obs, next_obs, actions, rewards, terminated, truncated, terminal_obs: Optional[torch.Tensor] = None):
        """OffPolicyAlgorithm._store_transition + ReplayBuffer.add for one VecEnv step (all device tensors)."""
        p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
        o, so = self._rows(obs, self.obs_dim) if obs is not None else (None, 0)  # None: add_obs() stored it already
        nx, sn = self._rows(next_obs, self.obs_dim)
        tb, stb = self._rows(terminal_obs, self.obs_dim) if terminal_obs is not None else (None, 0)
        actions = actions.reshape(self.n_envs, self.act_dim).to(torch.float32).contiguous()
        rewards = rewards.to(torch.float32).contiguous()
        terminated, truncated = terminated.to(torch.uint8).contiguous(), truncated.to(torch.uint8).contiguous()
        nat.check(nat.lib().usvgpu_replay_add(
            nat.device_index(self.device), self.n_envs, self.obs_dim, self.act_dim, self.buffer_size, self.pos, o, so, nx, sn,
            tb, stb, p(actions), p(rewards), p(terminated), p(truncated), p(self.observations),
            p(self.next_observations), p(self.actions), p(self.rewards), p(self.dones), p(self.timeouts),
            self._stream()))
        self.pos += 1
        if self.pos == self.buffer_size:
            self.full, self.pos = True, 0

    def add_obs(self, obs: torch.Tensor) -> None:
        """First half of a split add(): stores only the observation of slot ``pos`` (call before ``env.step``, which
        overwrites the observation tensor in place); finish with ``add(None, next_obs, ...)``."""
        o, so = self._rows(obs, self.obs_dim)
        p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
        nat.check(nat.lib().usvgpu_replay_add(
            nat.device_index(self.device), self.n_envs, self.obs_dim, self.act_dim, self.buffer_size, self.pos, o, so, None, 0,
            None, 0, None, None, None, None, p(self.observations), p(self.next_observations), p(self.actions),
            p(self.rewards), p(self.dones), p(self.timeouts), self._stream()))

    def sample(self, batch_size: int, generator: Optional[torch.Generator] = None, batch_inds=None, env_inds=None):
        """ReplayBuffer.sample: uniform over the filled slots and the envs (indices may be given, for tests)."""
        upper = self.buffer_size if self.full else self.pos
        if upper == 0:
            raise RuntimeError("sample() from an empty replay buffer")
        if batch_inds is None:
            batch_inds = torch.randint(0, upper, (batch_size,), device=self.device, generator=generator)
            env_inds = torch.randint(0, self.n_envs, (batch_size,), device=self.device, generator=generator)
        batch_inds = torch.as_tensor(batch_inds, dtype=torch.int64, device=self.device).contiguous()
        env_inds = torch.as_tensor(env_inds, dtype=torch.int64, device=self.device).contiguous()
        b = int(batch_inds.numel())
        e = lambda *shape: torch.empty(shape, dtype=torch.float32, device=self.device)  # noqa: E731
        out = ReplaySamples(e(b, self.obs_dim), e(b, self.act_dim), e(b, self.obs_dim), e(b, 1), e(b, 1))
        p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
        nat.check(nat.lib().usvgpu_replay_sample(
            nat.device_index(self.device), self.n_envs, self.obs_dim, self.act_dim, self.buffer_size, b, p(batch_inds),
            p(env_inds),
            p(self.observations), p(self.next_observations), p(self.actions), p(self.rewards), p(self.dones),
            p(self.timeouts), p(out.observations), p(out.next_observations), p(out.actions), p(out.rewards),
            p(out.dones), self._stream()))
        return out


def collect_transitions(env, actor, buf: DeviceReplayBuffer, last_obs: torch.Tensor, n_steps: int,
                        generator: Optional[torch.Generator] = None):
    """OffPolicyAlgorithm.collect_rollouts (train_freq = n_steps env steps) with device-resident tensors.

    env: UsvGpuVecEnv or UsvGpuVecFrameStack; actor: callable obs -> actions in the action box, or None for the
    uniform warm-up before ``learning_starts`` (SB3 samples the action space).  Returns the new last_obs.
    """
    venv = env
    while hasattr(venv, "venv"):  # UsvGpuVecFrameStack / UsvGpuVecVideoRecorder wrappers
        venv = venv.venv
    stacked = hasattr(env, "terminal_stack")
    low = torch.as_tensor(venv.action_space.low, device=last_obs.device, dtype=torch.float32)
    high = torch.as_tensor(venv.action_space.high, device=last_obs.device, dtype=torch.float32)
    for _ in range(n_steps):
        if actor is None:
            u = torch.rand((venv.num_envs, venv.act_dim), device=last_obs.device, generator=generator)
            actions = low + (high - low) * u
        else:
            with torch.no_grad():
                actions = torch.clamp(actor(last_obs), low, high)
        buf.add_obs(last_obs)  # before the step: the env overwrites its observation tensor in place
        new_obs, rewards, term, trunc = env.step_tensor(actions.contiguous())
        tobs = env.terminal_stack if stacked else venv.terminal_obs
        buf.add(None, new_obs, actions, rewards, term, trunc, tobs)
        last_obs = new_obs
    return last_obs
