"""The seats of the reference's callers (gym_usv_b200/sb3.py): SB3 VecEnv subclass, make_vec_env, gymnasium registry ids,
single-env gym.Env view; the info dict and the option-driven resets of the simple envs against goldens written by the
reference (tests/golden/simple_info.npz, oracle/gen_golden_info.py).

stable_baselines3 / gymnasium are not installed here: the MRO is driven with a stand-in that follows SB3's
``VecEnv.__init__(num_envs, observation_space, action_space)`` and declares the same abstract methods.
"""
import abc

import numpy as np
import pytest

from _helpers import inject_gpu, load_golden

INFO_KEYS = ("position", "velocity", "path_start", "path_end", "reward", "action0", "action1", "ye", "angle_to_target",
             "ye_reward", "angle_to_target_reward", "angle_action_reward", "delta_action_reward", "delta_action",
             "velocity_track_reward", "reference_velocity", "reward_velocity", "reference_velocity_error")


class StubVecEnv(abc.ABC):
    """stable_baselines3.common.vec_env.base_vec_env.VecEnv, reduced to its constructor contract and abstract set."""

    def __init__(self, num_envs, observation_space, action_space):
        self.num_envs, self.observation_space, self.action_space = num_envs, observation_space, action_space
        self.reset_infos = [{} for _ in range(num_envs)]
        self._seeds = [None for _ in range(num_envs)]
        self._options = [{} for _ in range(num_envs)]
        self.render_mode = self.get_attr("render_mode")[0]

    def _reset_options(self):
        self._options = [{} for _ in range(self.num_envs)]

    def set_options(self, options=None):
        options = options or {}
        self._options = [dict(options) for _ in range(self.num_envs)] if isinstance(options, dict) else list(options)

    @abc.abstractmethod
    def reset(self): ...
    @abc.abstractmethod
    def step_async(self, actions): ...
    @abc.abstractmethod
    def step_wait(self): ...
    @abc.abstractmethod
    def close(self): ...
    @abc.abstractmethod
    def get_attr(self, attr_name, indices=None): ...
    @abc.abstractmethod
    def set_attr(self, attr_name, value, indices=None): ...
    @abc.abstractmethod
    def env_method(self, method_name, *a, indices=None, **k): ...
    @abc.abstractmethod
    def env_is_wrapped(self, wrapper_class, indices=None): ...


# ------------------------------------------------------------------------------------------------ CPU
def test_sb3_class_mro_resolves_every_abstract_method():
    from gym_usv_b200 import UsvGpuVecEnv, sb3
    cls = sb3.sb3_vec_env_class(StubVecEnv)
    assert issubclass(cls, StubVecEnv) and issubclass(cls, UsvGpuVecEnv)
    assert cls.__mro__[1] is UsvGpuVecEnv  # the GPU implementation wins over the base's abstract declarations
    assert not getattr(cls, "__abstractmethods__", frozenset())
    for name in ("reset", "step_async", "step_wait", "close", "get_attr", "set_attr", "env_method", "env_is_wrapped"):
        assert not getattr(getattr(cls, name), "__isabstractmethod__", False), name


def test_sb3_class_constructor_calls_the_base_with_sb3_arguments(monkeypatch):
    from gym_usv_b200 import UsvGpuVecEnv, sb3, spaces

    def fake_init(self, num_envs, **kw):  # no CUDA here: stand in for the handle creation
        self.num_envs, self.render_mode = num_envs, kw.get("render_mode")
        self.observation_space, self.action_space = spaces.simple_observation_space(), spaces.simple_action_space()
        self.calls = []

    monkeypatch.setattr(UsvGpuVecEnv, "__init__", fake_init)
    monkeypatch.setattr(UsvGpuVecEnv, "get_attr", lambda self, name, indices=None: [getattr(self, name)] * self.num_envs)
    monkeypatch.setattr(UsvGpuVecEnv, "reset", lambda self, seed=None, options=None: self.calls.append(options) or "obs")
    env = sb3.sb3_vec_env_class(StubVecEnv)(8, render_mode="rgb_array")
    assert isinstance(env, StubVecEnv) and env.num_envs == 8 and env.render_mode == "rgb_array"
    assert len(env.reset_infos) == 8 and env.observation_space.shape == (143,)
    env.set_options({"place_obstacles_on_path": 2})
    assert env.reset() == "obs" and env.calls == [{"place_obstacles_on_path": 2}]
    assert env.reset() == "obs" and env.calls[-1] is None  # SB3 clears the options after a reset


def test_registry_mirrors_the_reference_ids_and_limits():
    from gym_usv_b200 import sb3
    # gym_usv/__init__.py:3-34 (usv-aitsmc-simple needs usv_libs_py: out of scope)
    assert sb3.REGISTRY == {"usv-asmc-v0": 0, "usv-pid-v0": 0, "usv-asmc-ye-int-v0": 0, "usv-asmc-ca-v0": 5000,
                            "usv-simple": 500, "usv-asmc-simple": 1000}
    assert sb3.register_envs() == [] or all(i in sb3.REGISTRY for i in sb3.register_envs(force=True))
    with pytest.raises(Exception):
        sb3.make_vec_env("usv-aitsmc-simple", n_envs=2)


def test_info_golden_holds_every_key_of_the_reference_dict():
    g = load_golden("simple_info.npz")
    for key in INFO_KEYS:
        assert f"info_asmc_info_{key}" in g.files and f"info_kin_info_{key}" in g.files
    assert int(g["path_reset_n_obs"]) == int(g["path_n_obs_without_option"]) + 3  # simple_env.py:276-288
    assert np.array_equal(g["exp_a_reset_ox"][:5], g["exp_obstacle_positions"][:, 0])


# ------------------------------------------------------------------------------------------------ GPU
@pytest.fixture(scope="module")
def usv():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import gym_usv_b200
    return gym_usv_b200


def _inject(env, g, prefix, keep_rows=()):
    st = env.get_state()
    rec = {k[len(prefix):]: g[k] for k in g.files if k.startswith(prefix + "reset_")}
    gg = {k: np.asarray(v)[None] for k, v in rec.items()}
    inject_gpu(st, 0, gg, 0, keep_rows=keep_rows)
    st["target0"][0] = g[prefix + "reset_target"]
    env.set_state(st)


@pytest.mark.gpu
@pytest.mark.parametrize("env_id,prefix", [("usv-asmc-simple", "info_asmc_"), ("usv-simple", "info_kin_")])
def test_step_info_matches_the_reference_dict(usv, env_id, prefix):
    import torch
    g = load_golden("simple_info.npz")
    env = usv.UsvGpuVecEnv(1, device=0, env_id=env_id, auto_reset=False)
    env.reset_tensor()
    _inject(env, g, prefix, keep_rows=(6, 7) if env_id == "usv-simple" else ())
    env.refresh_reset_obs()
    np.testing.assert_allclose(env.obs[0, :15].cpu().numpy(), g[prefix + "reset_obs"][:15], rtol=0, atol=2e-6)
    info0 = env.step_info()[0]
    for key in ("position", "velocity", "path_start", "path_end", "ye", "angle_to_target"):
        np.testing.assert_allclose(info0[key], g[f"{prefix}reset_info_{key}"], rtol=0, atol=1e-6, err_msg=key)
    acts = g[prefix + "actions"]
    for t in range(len(acts)):
        obs, rew, term, trunc = env.step_tensor(torch.tensor(acts[t:t + 1], device=env.device))
        assert bool(term[0]) == bool(g[prefix + "terminated"][t]) and bool(trunc[0]) == bool(g[prefix + "truncated"][t])
        np.testing.assert_allclose(obs[0].cpu().numpy(), g[prefix + "obs"][t], rtol=0, atol=2e-6)
        info = env.step_info()[0]
        assert set(INFO_KEYS) <= set(info) and info["left_thruster"] == 0 and info["right_thruster"] == 0
        for key in INFO_KEYS:
            tol = 2e-5 if key == "reward" else 1e-7  # the reward buffer is float32
            np.testing.assert_allclose(info[key], g[f"{prefix}info_{key}"][t], rtol=0, atol=tol, err_msg=f"{key} step {t}")
    env.close()


@pytest.mark.gpu
def test_custom_experiment_resets_and_rollout_match_the_reference(usv):
    import torch
    g = load_golden("simple_info.npz")
    exp = {k: g["exp_" + k] for k in ("obstacle_positions", "obstacle_radius", "path_start", "position")}
    exp["angle"] = float(g["exp_angle"])
    opts = {"run_custom_experiment": True, "experiment": exp}
    with pytest.raises(usv.UsvError):
        usv.UsvGpuVecEnv(1, device=0, env_id="usv-asmc-simple", options=opts)  # simple_env_asmc.py:10 takes no options
    env = usv.UsvGpuVecEnv(3, device=0, env_id="usv-simple", options=opts, max_episode_steps=0)  # the golden rollout has no TimeLimit
    env.reset()
    st = env.get_state()
    for i in range(3):  # the scene of simple_env.py:292-300 is installed in every env
        assert st["n_obs"][i] == 5
        np.testing.assert_array_equal(st["obstacles"][i, 0, :5], exp["obstacle_positions"][:, 0])
        np.testing.assert_array_equal(st["obstacles"][i, 2, :5], exp["obstacle_radius"])
        np.testing.assert_array_equal(st["dyn"][0:3, i], exp["position"])
        np.testing.assert_allclose(st["epi"][0:4, i], np.r_[g["exp_a_reset_path_start"], g["exp_a_reset_path_end"]], atol=1e-12)
    # the draws the experiment leaves random (velocity, max_action, v_ref, target) from the reference, then its rollout
    for prefix in ("exp_a_", "exp_b_"):
        st = env.get_state()
        st["dyn"][3:6, 0] = g[prefix + "reset_velocity"]
        st["epi"][4, 0], st["epi"][5, 0] = g[prefix + "reset_max_action"][[0, 2]]
        st["epi"][6, 0] = g[prefix + "reset_v_ref"]
        st["target0"][0] = g[prefix + "reset_target"]
        env.set_state(st)
        env.refresh_reset_obs()
        np.testing.assert_allclose(env.obs[0, :15].cpu().numpy(), g[prefix + "reset_obs"][:15], rtol=0, atol=2e-6)
        acts = g[prefix + "actions"]
        for t in range(len(acts)):
            a = np.repeat(acts[t:t + 1], 3, axis=0)
            obs, rew, dones, infos = env.step(a)
            last = t == len(acts) - 1 and (g[prefix + "terminated"][t] or g[prefix + "truncated"][t])
            row = infos[0]["terminal_observation"] if last else obs[0]
            np.testing.assert_allclose(row, g[prefix + "obs"][t], rtol=0, atol=2e-6, err_msg=f"{prefix} step {t}")
            np.testing.assert_allclose(rew[0], g[prefix + "reward"][t], rtol=0, atol=2e-5)
            assert bool(dones[0]) == bool(g[prefix + "terminated"][t] or g[prefix + "truncated"][t])
            if dones[0]:
                break
        # a finished env was reset through the host: the scene is back (every reset re-installs it, :292)
        env.reset()
        st = env.get_state()
        np.testing.assert_array_equal(st["dyn"][0:3, 0], exp["position"])
        assert st["n_obs"][0] == 5 and st["steps"][0] == 0
    # a TimeLimit with the experiment installed: the in-step reset of finished envs runs through the host edits
    env2 = usv.UsvGpuVecEnv(2, device=0, env_id="usv-simple", options=opts, max_episode_steps=5)
    env2.reset()
    a = np.tile(np.array([[0.3, 0.0]], np.float32), (2, 1))
    for t in range(5):
        obs, rew, dones, infos = env2.step(a)
    assert dones.all() and all("terminal_observation" in i and i["TimeLimit.truncated"] for i in infos)
    st = env2.get_state()
    np.testing.assert_array_equal(st["dyn"][0:3, 1], exp["position"])
    assert (st["steps"] == 0).all() and (st["n_obs"] == 5).all()
    env.close()
    env2.close()


@pytest.mark.gpu
def test_place_obstacles_on_path_option(usv):
    g = load_golden("simple_info.npz")
    base = usv.UsvGpuVecEnv(64, device=0, env_id="usv-simple", seed=7)
    base.reset()
    n0 = base.get_state(["n_obs"])["n_obs"].copy()
    env = usv.UsvGpuVecEnv(64, device=0, env_id="usv-simple", seed=7)
    obs = env.reset(options={"place_obstacles_on_path": 3}).copy()
    st = env.get_state()
    assert np.array_equal(st["n_obs"], n0 + 3)  # simple_env.py:286-288: appended after the rejection step
    for i in range(64):
        k = int(n0[i])
        ps, pe = st["epi"][0:2, i], st["epi"][2:4, i]
        d = (pe - ps) / np.hypot(*(pe - ps))
        rel = np.stack([st["obstacles"][i, 0, k:k + 3] - ps[0], st["obstacles"][i, 1, k:k + 3] - ps[1]], axis=1)
        along, across = rel @ d, rel @ np.array([-d[1], d[0]])
        assert np.all(np.abs(across) < 6.0) and np.all(along > -6.0) and np.all(along < 26.0)  # N(path point, 1), mag in [0, 20]
        assert np.all((st["obstacles"][i, 2, k:k + 3] >= 0.15) & (st["obstacles"][i, 2, k:k + 3] <= 0.5))
    # 'usv-asmc-simple' drops reset options like the reference (simple_env_asmc.py:14-16)
    e2 = usv.UsvGpuVecEnv(8, device=0, env_id="usv-asmc-simple", seed=7)
    e2.reset()
    m0 = e2.get_state(["n_obs"])["n_obs"].copy()
    e2.reset(seed=7, options={"place_obstacles_on_path": 3})
    assert np.array_equal(e2.get_state(["n_obs"])["n_obs"], m0)
    # the reference's own reset with the option: its record, installed here, gives its reset observation
    one = usv.UsvGpuVecEnv(1, device=0, env_id="usv-simple", auto_reset=False)
    one.reset_tensor()
    _inject(one, g, "path_", keep_rows=(6, 7))
    one.refresh_reset_obs()
    np.testing.assert_allclose(one.obs[0, :15].cpu().numpy(), g["path_reset_obs"][:15], rtol=0, atol=2e-6)
    with pytest.raises(usv.UsvError):
        env.reset(options={"place_obstacles_on_path": 40})
    for e in (base, env, e2, one):
        e.close()
    assert obs.shape == (64, 143)


@pytest.mark.gpu
def test_sb3_seat_and_gym_view_on_the_device(usv):
    from gym_usv_b200 import sb3
    cls = sb3.sb3_vec_env_class(StubVecEnv)
    env = cls(32, device=0, env_id="usv-asmc-simple", seed=2)
    assert isinstance(env, StubVecEnv) and env.num_envs == 32 and len(env.reset_infos) == 32
    obs = env.reset()
    assert obs.shape == (32, 143) and obs.dtype == np.float32
    a = np.tile(np.array([[0.6, 0.1]], np.float32), (32, 1))
    obs, rew, dones, infos = env.step(a)
    assert obs.shape == (32, 143) and rew.shape == (32,) and dones.shape == (32,) and len(infos) == 32
    env.close()
    vec = sb3.make_vec_env("usv-simple", n_envs=4, seed=5, env_kwargs={"render_mode": "rgb_array"})
    assert vec.num_envs == 4 and vec.env_id == "usv-simple" and vec.render_mode == "rgb_array"
    vec.close()
    # gym.Env view: reset(seed, options) -> (obs, info), 5-tuple step; same numbers as env 0 of the batched env
    one = sb3.UsvGpuEnv("usv-asmc-simple", seed=2)
    ref = usv.UsvGpuVecEnv(1, device=0, env_id="usv-asmc-simple", seed=2, auto_reset=False)
    o, info = one.reset()
    assert np.array_equal(o, ref.reset()[0]) and info["reward"] == -1.0 and set(info) >= {"position", "ye", "angle_to_target"}
    for t in range(20):
        o, r, term, trunc, info = one.step(np.array([0.7, -0.2], np.float32))
        ob, rb, db, _ = ref.step(np.array([[0.7, -0.2]], np.float32))
        assert np.array_equal(o, ob[0]) and r == float(rb[0]) and (term or trunc) == bool(db[0])
        assert set(INFO_KEYS) <= set(info)
        if term or trunc:
            break
    old = sb3.UsvGpuEnv("usv-asmc-v0", seed=1, legacy_api=True)  # usv_asmc_env.py:255,300: old-gym tuples
    assert old.reset().shape == (6,)
    step = old.step(np.array([0.1], np.float32))
    assert len(step) == 4 and step[0].shape == (6,)
    for e in (one, ref, old):
        e.close()
