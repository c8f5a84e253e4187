"""The seats the reference's callers expect: an SB3 ``VecEnv`` subclass, ``make_vec_env``, the gymnasium registry and
a single-env ``gym.Env`` view -- all over the same CUDA library (no CPU path anywhere).

What the reference's scripts do (train_test/sb3_train_vec.py:67-72, train_test/sb3_train.py:49, tools/profile_env.py:5-7):

    env = make_vec_env(env_id, n_envs=4, env_kwargs={"render_mode": "rgb_array"})   # -> DummyVecEnv of gym.make(id)
    env = VecVideoRecorder(env, ...); env = VecFrameStack(env, n_stack=5); SAC("MlpPolicy", env, ...)
    env = gym.make('usv-simple'); env.reset(); env.step(np.zeros(2))

and what this module gives them:

    make_vec_env(env_id, n_envs, seed=None, env_kwargs=None)   same signature as SB3's; returns the GPU VecEnv
    UsvGpuSb3VecEnv       class UsvGpuSb3VecEnv(UsvGpuVecEnv, stable_baselines3.common.vec_env.VecEnv): exists when SB3
                          imports (``BaseAlgorithm._wrap_env`` then takes it as a VecEnv instead of wrapping it in a
                          DummyVecEnv); ``sb3_vec_env_class(base)`` builds the same class over any base with SB3's
                          ``VecEnv.__init__(num_envs, observation_space, action_space)`` contract
    register_envs()       the six ids of gym_usv/__init__.py:3-34 with their step limits, ``entry_point`` = the single-env
                          view, ``vector_entry_point`` = the batched env (gymnasium >= 0.29 ``gym.make_vec``)
    UsvGpuEnv             ``reset(seed=None, options=None) -> (obs, info)``, ``step(a) -> (obs, r, terminated, truncated,
                          info)`` (gym_usv/envs/simple_env.py:228,310,346) over a one-env handle; the legacy ids keep the
                          old-gym 4-tuple of usv_asmc_env.py:255 when ``legacy_api=True``

stable_baselines3 and gymnasium are absent from the build image: everything here imports without them, and the tests
drive the MRO with a stand-in base that follows SB3's ``VecEnv.__init__``.
"""
from __future__ import annotations

from typing import Any, Optional

import numpy as np

from . import _native as nat
from .vec_env import ENV_KINDS, LEGACY_IDS, MAX_EPISODE_STEPS, UsvGpuVecEnv

# gym_usv/__init__.py:3-34 (usv-aitsmc-simple, :36-40, runs a controller that exists only in usv_libs_py: not built)
REGISTRY = {env_id: MAX_EPISODE_STEPS[env_id] for env_id in ENV_KINDS}


def _import_sb3_vec_env():
    try:
        from stable_baselines3.common.vec_env import VecEnv  # type: ignore
        return VecEnv
    except Exception:  # noqa: BLE001
        return None


def _import_gymnasium():
    try:
        import gymnasium  # type: ignore
        return gymnasium
    except Exception:  # noqa: BLE001
        return None


def sb3_vec_env_class(base):
    """``class UsvGpuSb3VecEnv(UsvGpuVecEnv, base)`` for an SB3-style ``VecEnv`` base.

    UsvGpuVecEnv comes first in the MRO, so every abstract method of ``VecEnv`` (reset, step_async, step_wait, close,
    get_attr, set_attr, env_method, env_is_wrapped) resolves to the GPU implementation; the base contributes
    ``isinstance(env, VecEnv)``, ``step()``, ``unwrapped``, ``getattr_depth_check`` and the wrapper plumbing."""

    class UsvGpuSb3VecEnv(UsvGpuVecEnv, base):  # type: ignore[misc, valid-type]
        def __init__(self, num_envs: int, *args, **kwargs):
            UsvGpuVecEnv.__init__(self, num_envs, *args, **kwargs)
            render_mode = self.render_mode
            # SB3: VecEnv.__init__(num_envs, observation_space, action_space) -- sets reset_infos / _seeds / _options
            # and reads render_mode back through get_attr()
            base.__init__(self, self.num_envs, self.observation_space, self.action_space)
            self.render_mode = render_mode

        def reset(self):  # SB3 VecEnv.reset(): seeds / options come from seed() / set_options()
            opts = getattr(self, "_options", None)
            options = opts[0] if opts else None
            obs = UsvGpuVecEnv.reset(self, options=options or None)
            if hasattr(self, "_reset_options"):
                self._reset_options()
            return obs

        def set_options(self, options=None):  # SB3 keeps them in self._options until the next reset()
            if hasattr(base, "set_options"):
                base.set_options(self, options)
            else:
                UsvGpuVecEnv.set_options(self, options)

        def step(self, actions):
            self.step_async(actions)
            return self.step_wait()

    UsvGpuSb3VecEnv.__qualname__ = "UsvGpuSb3VecEnv"
    return UsvGpuSb3VecEnv


_SB3_BASE = _import_sb3_vec_env()
UsvGpuSb3VecEnv = sb3_vec_env_class(_SB3_BASE) if _SB3_BASE is not None else None


def make_vec_env(env_id: str, n_envs: int = 1, seed: Optional[int] = None, start_index: int = 0, monitor_dir=None,
                 wrapper_class=None, env_kwargs: Optional[dict] = None, vec_env_cls=None, vec_env_kwargs: Optional[dict] = None,
                 monitor_kwargs=None, wrapper_kwargs=None):
    """Drop-in for ``stable_baselines3.common.env_util.make_vec_env`` (train_test/sb3_train_vec.py:67-68): the n_envs
    copies of ``env_id`` as ONE batched GPU env.  ``vec_env_cls`` (DummyVecEnv / SubprocVecEnv in the reference) is
    ignored -- there are no per-env processes; Monitor's episode statistics come from the device
    (``infos[i]["episode"]``), so ``monitor_dir`` / ``monitor_kwargs`` are accepted and unused.  ``wrapper_class`` is a
    per-env gym.Wrapper in SB3 and has nothing to wrap here: passing one raises."""
    if env_id not in ENV_KINDS:
        raise nat.UsvError(f"unknown env id '{env_id}' (have: {sorted(ENV_KINDS)})")
    if wrapper_class is not None:
        raise nat.UsvError("make_vec_env: wrapper_class wraps single gym.Env objects; the GPU env has none to wrap")
    kw = dict(env_kwargs or {})
    kw.update(vec_env_kwargs or {})
    kw.setdefault("seed", 0 if seed is None else int(seed))
    kw.setdefault("global_env_offset", int(start_index))
    cls = UsvGpuSb3VecEnv if UsvGpuSb3VecEnv is not None else UsvGpuVecEnv
    return cls(int(n_envs), env_id=env_id, **kw)


class UsvGpuEnv:
    """One env behind the gymnasium ``Env`` API (gym_usv/envs/simple_env.py:228,310,346): a one-env UsvGpuVecEnv with the
    in-kernel auto-reset off -- an episode ends with terminated / truncated and the caller resets, as ``gym.make`` users
    (tools/profile_env.py:5-7, train_test/sb3_train.py:49) expect.  ``info`` carries the reference's keys
    (simple_env.py:102-115, 189-199) for the simple envs."""

    metadata = UsvGpuVecEnv.metadata

    def __init__(self, env_id: str = "usv-asmc-simple", render_mode: Optional[str] = None, device=0, seed: int = 0,
                 legacy_api: bool = False, options: Optional[dict] = None, **kwargs):
        kwargs.pop("auto_reset", None)
        self.env_id = env_id
        self.legacy_api = bool(legacy_api) and env_id in LEGACY_IDS
        self.vec = UsvGpuVecEnv(1, device=device, seed=seed, env_id=env_id, render_mode=render_mode, auto_reset=False,
                                options=options, **kwargs)
        self.observation_space, self.action_space = self.vec.observation_space, self.vec.action_space
        self.render_mode = render_mode
        self.spec = None
        self.np_random = None
        self._episode = 0

    @property
    def unwrapped(self):
        return self

    def reset(self, *, seed: Optional[int] = None, options: Optional[dict] = None):
        if seed is not None:
            self.vec.seed(int(seed))
        obs = self.vec.reset(options=options)
        self._episode += 1
        obs0 = np.array(obs[0], copy=True)
        if self.legacy_api:
            return obs0  # usv_asmc_env.py:300 returns the state only
        return obs0, self._info(reward=-1.0)  # simple_env.py:305 _get_info(-1, zeros)

    def step(self, action):
        a = np.asarray(action, dtype=np.float32).reshape(1, self.vec.act_dim)
        obs, rew, dones, infos = self.vec.step(a)
        term, trunc = bool(self.vec._host[self.vec._slot]["term"][0]), bool(self.vec._host[self.vec._slot]["trunc"][0])
        obs0, r = np.array(obs[0], copy=True), float(rew[0])
        info = self._info(reward=r)
        info.update({k: v for k, v in infos[0].items() if k == "episode"})
        if self.legacy_api:
            return obs0, r, term or trunc, info  # usv_asmc_env.py:255
        return obs0, r, term, trunc, info

    def _info(self, reward: float) -> dict:
        if self.vec.legacy or self.vec.ca:
            return {}
        info = self.vec.step_info(indices=[0])[0]
        info["reward"] = reward
        return info

    def render(self):
        if self.render_mode != "rgb_array":
            return None
        return self.vec.render_tensor(indices=[0])[0].cpu().numpy()

    def close(self):
        self.vec.close()


def _make_single(env_id):
    def factory(render_mode: Optional[str] = None, **kwargs):
        return UsvGpuEnv(env_id, render_mode=render_mode, **kwargs)
    return factory


def _make_vector(env_id):
    def factory(num_envs: int = 1, **kwargs):
        return make_vec_env(env_id, n_envs=num_envs, env_kwargs=kwargs)
    return factory


def register_envs(namespace: Optional[str] = None, force: bool = False) -> list:
    """Registers the ids of gym_usv/__init__.py:3-34 with gymnasium (same ids, same ``max_episode_steps``), each with
    ``entry_point`` = UsvGpuEnv and ``vector_entry_point`` = the batched GPU env.  ``namespace="b200"`` registers
    ``b200/usv-asmc-simple`` etc. instead, to live beside an imported ``gym_usv``.  Ids that are already registered are
    left alone unless ``force``.  Returns the ids registered; [] when gymnasium is not installed."""
    gymnasium = _import_gymnasium()
    if gymnasium is None or not hasattr(gymnasium, "registry") or not hasattr(gymnasium, "register"):
        return []  # not installed (or an import-only stand-in, as the oracle harness installs)
    done = []
    for env_id, limit in REGISTRY.items():
        full = f"{namespace}/{env_id}" if namespace else env_id
        if full in gymnasium.registry and not force:
            continue
        kw = dict(id=full, entry_point=_make_single(env_id), max_episode_steps=(limit or None))
        try:
            gymnasium.register(vector_entry_point=_make_vector(env_id), **kw)
        except TypeError:  # gymnasium < 0.29 has no vector_entry_point
            gymnasium.register(**kw)
        done.append(full)
    return done
