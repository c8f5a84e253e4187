"""Import the UNMODIFIED reference (/root/reference) in this container.

TEST INFRASTRUCTURE ONLY.  Nothing under ``gym_usv_b200/`` may import this.
It exists so that ``oracle/gen_golden.py`` can run the reference's own code and
write the golden vectors under ``tests/golden/``, and so that
``tests/test_oracle_vs_reference.py`` can pin the C restatement against the
live reference when ``/root/reference`` is present (it is not on the GPU box).

What it does (SURVEY.md section 8c recipe):
  * ``np.math = math``  -- numpy >= 2 dropped the alias the reference uses at
    gym_usv/control/usv_asmc.py:72 and gym_usv/envs/usv_asmc_env.py:234,281.
  * import-only stub modules for packages that are absent here and carry no
    arithmetic on the path: ``gymnasium`` (Env.reset seeds ``np_random`` with
    ``np.random.default_rng(seed)``, which is what gymnasium's
    ``seeding.np_random`` does: PCG64(SeedSequence(seed))), ``gym``, ``pygame``,
    ``matplotlib.pyplot``, ``usv_libs_py``.
The reference sources are imported from where they lie; nothing is copied.
"""
from __future__ import annotations

import math
import os
import sys
import types

import numpy as np

REFERENCE_ROOT = os.environ.get("USV_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "gym_usv", "control", "usv_asmc.py"))


class _Box:
    def __init__(self, low, high, shape=None, dtype=np.float32):
        self.low, self.high, self.dtype = low, high, dtype
        self.shape = tuple(shape) if shape is not None else np.shape(low)


class _GymnasiumEnv:
    """gymnasium.Env stand-in: only the seeding behaviour of reset()."""
    np_random = None

    def reset(self, seed=None, options=None):
        if seed is not None or self.np_random is None:
            self.np_random = np.random.default_rng(seed)


class _LegacyGymEnv:
    pass


def _mod(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


class _Anything:
    """Attribute sink for pygame / usv_libs_py symbols touched at import or __init__."""

    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return _Anything()

    def __getattr__(self, name):
        return _Anything()


_installed = False


def install_stubs() -> None:
    global _installed
    if _installed:
        return
    if not hasattr(np, "math"):
        np.math = math  # usv_asmc.py:72 uses np.math.asin
    spaces = _mod("gymnasium.spaces", Box=_Box)
    _mod("gymnasium", Env=_GymnasiumEnv, spaces=spaces, register=lambda *a, **k: None)
    gspaces = _mod("gym.spaces", Box=_Box)
    gseeding = _mod("gym.utils.seeding")
    gutils = _mod("gym.utils", seeding=gseeding)
    gerror = _mod("gym.error")
    _mod("gym", Env=_LegacyGymEnv, spaces=gspaces, utils=gutils, error=gerror)
    pg = _mod("pygame")
    pg.__getattr__ = lambda name: _Anything()  # type: ignore[attr-defined]
    _mod("matplotlib")
    _mod("matplotlib.pyplot")
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    ctrl = _mod("usv_libs_py.controller", ASMC=_Anything, ASMCSetpoint=_Anything, AITSMC=_Anything)
    model = _mod("usv_libs_py.model", DynamicModel=_Anything)
    utils = _mod("usv_libs_py.utils")
    utils.__getattr__ = lambda name: _Anything()  # type: ignore[attr-defined]
    _mod("usv_libs_py", controller=ctrl, model=model, utils=utils)
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    _installed = True


def load_reference():
    """Returns a namespace with the reference classes on the hot path."""
    if not reference_available():
        raise RuntimeError(f"reference not found under {REFERENCE_ROOT}")
    install_stubs()
    from gym_usv.control.usv_asmc import UsvAsmc
    from gym_usv.envs.simple_env import UsvSimpleEnv
    from gym_usv.envs.simple_env_asmc import UsvSimpleASMCEnv
    from gym_usv.envs.usv_asmc_ca_env import UsvAsmcCaEnv
    from gym_usv.envs.usv_asmc_env import UsvAsmcEnv
    from gym_usv.envs.usv_asmc_ye_int_env import UsvAsmcYeIntEnv
    from gym_usv.envs.usv_pid_env import UsvPidEnv

    return types.SimpleNamespace(
        UsvAsmc=UsvAsmc,
        UsvSimpleEnv=UsvSimpleEnv,
        UsvSimpleASMCEnv=UsvSimpleASMCEnv,
        UsvAsmcCaEnv=UsvAsmcCaEnv,
        UsvAsmcEnv=UsvAsmcEnv,
        UsvAsmcYeIntEnv=UsvAsmcYeIntEnv,
        UsvPidEnv=UsvPidEnv,
    )


class TimeLimit:
    """gymnasium.wrappers.TimeLimit behaviour (gym_usv/__init__.py:30-34 registers
    usv-asmc-simple with max_episode_steps=1000): truncated |= elapsed >= limit."""

    def __init__(self, env, max_episode_steps):
        self.env, self.max_episode_steps, self.elapsed = env, max_episode_steps, 0

    def reset(self, **kw):
        self.elapsed = 0
        return self.env.reset(**kw)

    def step(self, action):
        obs, rew, term, trunc, info = self.env.step(action)
        self.elapsed += 1
        if self.elapsed >= self.max_episode_steps:
            trunc = True
        return obs, rew, term, trunc, info
