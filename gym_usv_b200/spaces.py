"""Observation / action spaces of the reference envs.

When gymnasium is importable the very same ``gymnasium.spaces.Box`` objects the reference builds
are returned (gym_usv/envs/simple_env.py:27,30; gym_usv/envs/usv_asmc_env.py:89-96); otherwise a
minimal stand-in with the attributes SB3 reads (low, high, shape, dtype, sample, contains).
"""
from __future__ import annotations

import numpy as np

try:  # pragma: no cover - gymnasium is absent in the build container
    from gymnasium.spaces import Box as _GymBox
except Exception:  # noqa: BLE001
    _GymBox = None


class Box:
    """Stand-in for gymnasium.spaces.Box (used only when gymnasium is not installed)."""

    def __init__(self, low, high, shape=None, dtype=np.float32):
        self.dtype = np.dtype(dtype)
        if shape is None:
            shape = np.shape(low)
        self.shape = tuple(shape)
        self.low = np.broadcast_to(np.asarray(low, dtype=self.dtype), self.shape).copy()
        self.high = np.broadcast_to(np.asarray(high, dtype=self.dtype), self.shape).copy()
        self._rng = np.random.default_rng()

    def seed(self, seed=None):
        self._rng = np.random.default_rng(seed)
        return [seed]

    def sample(self):
        return self._rng.uniform(self.low, self.high).astype(self.dtype)

    def contains(self, x):
        x = np.asarray(x)
        return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))

    def __repr__(self):
        return f"Box({self.low.min()}, {self.high.max()}, {self.shape}, {self.dtype})"


def make_box(low, high, shape=None, dtype=np.float32):
    if _GymBox is not None:
        return _GymBox(low=np.asarray(low, dtype=dtype) if np.ndim(low) else low,
                       high=np.asarray(high, dtype=dtype) if np.ndim(high) else high, shape=shape, dtype=dtype)
    return Box(low, high, shape=shape, dtype=dtype)


def simple_observation_space():
    """simple_env.py:27  Box(-1, 1, (15 + 128,), float32)  (not enforced/clipped by the reference)."""
    return make_box(-1.0, 1.0, shape=(143,), dtype=np.float32)


def simple_action_space():
    """simple_env.py:30  Box([0.2, -1], [1, 1], (2,), float32)."""
    return make_box(np.array([0.2, -1.0]), np.array([1.0, 1.0]), shape=(2,), dtype=np.float32)


def legacy_observation_space():
    """usv_asmc_env.py:89-96."""
    low = np.array([-1.5, -1.5, -1.0, -10.0, -np.pi, -np.pi / 2], dtype=np.float32)
    return make_box(low, -low, shape=(6,), dtype=np.float32)


def legacy_action_space():
    """usv_asmc_env.py:92."""
    return make_box(-np.pi / 2, np.pi / 2, shape=(1,), dtype=np.float32)


def ca_observation_space():
    """usv_asmc_ca_env.py:99-113  Box(-1, 1, (7 + 16,), float32)."""
    return make_box(-1.0, 1.0, shape=(23,), dtype=np.float32)


def ca_action_space():
    """usv_asmc_ca_env.py:106-110  Box([-1, -1], [1, 1], float32)."""
    return make_box(-1.0, 1.0, shape=(2,), dtype=np.float32)
