"""Golden vectors for the per-step info dict and for the option-driven resets of the simple envs, written by the
UNMODIFIED reference imported from /root/reference (oracle/ref_harness.py).  TEST INFRASTRUCTURE ONLY.

    python oracle/gen_golden_info.py      ->  tests/golden/simple_info.npz

Content:
  info_<env>_*     UsvSimpleASMCEnv ('asmc') and UsvSimpleEnv ('kin'): reset(seed) record, T actions, and per step every
                   numeric entry of the info dict (simple_env.py:102-115 | 189-199), up to the first done
  exp_*            UsvSimpleEnv(options={'run_custom_experiment': True, 'experiment': ...}) (simple_env.py:10,292-300):
                   the experiment, two resets (the scene is re-installed by every reset), a rollout
  path_*           UsvSimpleEnv.reset(seed, options={'place_obstacles_on_path': 3}) (simple_env.py:276-288): reset record
                   and observation (the extra obstacles lie within a few metres of the path)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_harness as rh  # noqa: E402
from gen_golden import _reset_record  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")
INFO_KEYS = ("position", "velocity", "path_start", "path_end", "reward", "action0", "action1", "ye", "angle_to_target",
             "ye_reward", "angle_to_target_reward", "angle_action_reward", "delta_action_reward", "delta_action",
             "velocity_track_reward", "reference_velocity", "reward_velocity", "reference_velocity_error")


def rollout(env, acts, out, prefix, seed, reset_kw=None):
    o, info0 = env.reset(seed=seed, **(reset_kw or {}))
    rec = _reset_record(env, o)
    for k, v in rec.items():
        out[f"{prefix}reset_{k}"] = np.asarray(v)
    for key in INFO_KEYS:  # reset() returns _get_info(-1, zeros) only (simple_env.py:305): no reward_info keys
        if key in info0:
            out[f"{prefix}reset_info_{key}"] = np.asarray(info0[key], dtype=np.float64)
    rows = {key: [] for key in INFO_KEYS}
    obs, rew, term, trunc = [], [], [], []
    for a in acts:
        o, r, te, tr, info = env.step(a)
        obs.append(o), rew.append(r), term.append(te), trunc.append(tr)
        for key in INFO_KEYS:
            rows[key].append(np.asarray(info[key], dtype=np.float64))
        if te or tr:
            break
    out[prefix + "actions"] = np.asarray(acts[:len(obs)], dtype=np.float32)
    out[prefix + "obs"], out[prefix + "reward"] = np.asarray(obs, np.float32), np.asarray(rew, np.float64)
    out[prefix + "terminated"], out[prefix + "truncated"] = np.asarray(term, np.uint8), np.asarray(trunc, np.uint8)
    for key in INFO_KEYS:
        out[f"{prefix}info_{key}"] = np.asarray(rows[key])
    return len(obs)


def main():
    R = rh.load_reference()
    rng = np.random.default_rng(11)
    out = {}
    acts = rng.uniform([0.2, -1.0], [1.0, 1.0], size=(120, 2)).astype(np.float32)
    n = rollout(R.UsvSimpleASMCEnv(), acts, out, "info_asmc_", seed=3)
    print("asmc info steps", n)
    n = rollout(R.UsvSimpleEnv(), acts, out, "info_kin_", seed=4)
    print("kin info steps", n)
    # custom experiment: the two hand-made layouts of tools/test_env.py:8-44 are for the CA env; this one is a scene of the
    # same kind for UsvSimpleEnv (obstacles on and beside a straight path)
    exp = dict(obstacle_positions=np.array([[12.0, 10.5], [14.0, 9.2], [16.5, 10.8], [11.0, 13.0], [15.0, 12.5]]),
               obstacle_radius=np.array([0.3, 0.45, 0.2, 0.5, 0.35]), path_start=np.array([9.0, 10.0]), angle=0.15,
               position=np.array([9.5, 9.8, 0.4]))
    for k, v in exp.items():
        out["exp_" + k] = np.asarray(v, dtype=np.float64)
    env = R.UsvSimpleEnv(options={"run_custom_experiment": True, "experiment": exp})
    n = rollout(env, acts, out, "exp_a_", seed=5)
    n2 = rollout(env, acts[::-1].copy(), out, "exp_b_", seed=None)  # a second reset of the same env object
    print("experiment steps", n, n2)
    env = R.UsvSimpleEnv()
    n = rollout(env, acts[:10], out, "path_", seed=6, reset_kw=dict(options={"place_obstacles_on_path": 3}))
    env2 = R.UsvSimpleEnv()
    o2, _ = env2.reset(seed=6)
    out["path_n_obs_without_option"] = np.asarray(env2.obstacle_n)
    print("path option: obstacles", int(out["path_reset_n_obs"]), "without:", int(env2.obstacle_n))
    np.savez_compressed(os.path.join(OUT, "simple_info.npz"), **out)


if __name__ == "__main__":
    main()
