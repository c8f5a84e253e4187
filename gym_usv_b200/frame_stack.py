"""UsvGpuVecFrameStack -- stable_baselines3 ``VecFrameStack`` for the device-resident VecEnv.

The reference's training scripts wrap the env as ``VecFrameStack(env, n_stack=5)`` (train_test/sb3_train_vec.py:70,
train_test/sb3_train.py:50-51).  SB3's ``StackedObservations.update`` rolls a ``[N, n_stack * obs_dim]`` array every
step (4 frames read + 5 written per env).  Here the frames never move: ``usvgpu_frame_stack_push`` keeps a ring of
``2 * n_stack`` slots per env and stores every new frame twice, so the ``n_stack`` latest frames are always one
contiguous run of the env's row; the stacked observation handed to the policy is a strided *view* of the ring
(row stride ``2 * n_stack * obs_dim`` floats, unit inner stride -- a valid GEMM operand, no copy).

Semantics follow SB3 (oldest frame first; zeros before the first frames of an episode; for finished envs the stacked
``terminal_observation`` = previous ``n_stack - 1`` frames + the env's terminal observation).
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _native as nat


class UsvGpuVecFrameStack:
    def __init__(self, venv, n_stack: int = 5):
        if n_stack < 1:
            raise ValueError("n_stack must be >= 1")
        self.venv, self.n_stack = venv, int(n_stack)
        self.num_envs, self.device = venv.num_envs, venv.device
        self.obs_dim = venv.obs_dim
        n, s, d = self.num_envs, self.n_stack, self.obs_dim
        self.ring = torch.zeros((n, 2 * s, d), dtype=torch.float32, device=self.device)
        self.terminal_stack = torch.zeros((n, s * d), dtype=torch.float32, device=self.device)
        self._k = 0  # frames pushed since the last reset()
        self._lib = nat.lib()
        self.action_space = venv.action_space
        self.observation_space = self._stacked_space(venv.observation_space)

    def _stacked_space(self, space):
        low = np.tile(np.asarray(space.low, dtype=np.float32), self.n_stack)
        high = np.tile(np.asarray(space.high, dtype=np.float32), self.n_stack)
        from . import spaces  # noqa: PLC0415
        box = spaces._GymBox if spaces._GymBox is not None else spaces.Box
        return box(low=low, high=high, dtype=np.float32)

    # ------------------------------------------------------------------ device face
    def _view(self, pos: int) -> torch.Tensor:
        n, s, d = self.num_envs, self.n_stack, self.obs_dim
        return torch.as_strided(self.ring, (n, s * d), (2 * s * d, 1), storage_offset=(pos + 1) * d)

    def _push(self, obs, term, trunc, tobs):
        pos = self._k % self.n_stack
        p = lambda t: None if t is None else C.c_void_p(t.data_ptr())  # noqa: E731
        nat.check(self._lib.usvgpu_frame_stack_push(
            nat.device_index(self.device), self.num_envs, self.obs_dim, self.n_stack, pos, p(self.ring), p(obs), p(term),
            p(trunc), p(tobs), p(self.terminal_stack) if tobs is not None else None,
            C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)))
        self._k += 1
        return self._view(pos)

    def reset_tensor(self) -> torch.Tensor:
        obs = self.venv.reset_tensor()
        self.ring.zero_()
        self._k = 0
        return self._push(obs, None, None, None)

    def step_tensor(self, actions: torch.Tensor):
        """-> (stacked_obs view [N, n_stack * obs_dim], reward, terminated, truncated); the view is valid until the
        next step (the ring slot it ends at is overwritten n_stack steps later, its oldest slot on the next step)."""
        obs, rew, term, trunc = self.venv.step_tensor(actions)
        tobs = self.venv.terminal_obs if self.venv.cfg.auto_reset else None
        return self._push(obs, term, trunc, tobs), rew, term, trunc

    # ------------------------------------------------------------------ host face (numpy, SB3 call shape)
    def reset(self) -> np.ndarray:
        return self.reset_tensor().cpu().numpy()

    def step(self, actions):
        a = torch.as_tensor(np.asarray(actions, dtype=np.float32).reshape(self.num_envs, self.venv.act_dim),
                            device=self.device)
        stacked, rew, term, trunc = self.step_tensor(a)
        term_h, trunc_h = term.cpu().numpy().astype(bool), trunc.cpu().numpy().astype(bool)
        dones = term_h | trunc_h
        infos = [{} for _ in range(self.num_envs)]
        idx = np.flatnonzero(dones)
        if idx.size:
            ts = self.terminal_stack[torch.as_tensor(idx, device=self.device)].cpu().numpy()
            for k, i in enumerate(idx):
                infos[int(i)] = {"TimeLimit.truncated": bool(trunc_h[i] and not term_h[i])}
                if self.venv.cfg.auto_reset:
                    infos[int(i)]["terminal_observation"] = ts[k]
        return stacked.cpu().numpy(), rew.cpu().numpy(), dones, infos

    def close(self):
        self.venv.close()
