"""Shared helpers for the parity tests (test infrastructure)."""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name))


def action_stream(n, seed=1):
    """SURVEY 8d config 1a/1b action stream (same as oracle/gen_golden.py)."""
    return np.random.default_rng(seed).uniform([0.2, -1.0], [1.0, 1.0], size=(n, 2)).astype(np.float32)


def golden_reset_maps(g):
    """{env: {step: reset record index}} for a simple_asmc_env golden file."""
    out = {}
    for k, (e, t) in enumerate(zip(g["reset_env"], g["reset_step"])):
        out.setdefault(int(e), {})[int(t)] = k
    return out


def inject_oracle(env_arr, i, g, k):
    """Put golden reset record k into oracle env i (numpy structured array, usv_oracle.ENV_DTYPE)."""
    import usv_oracle as uo
    env_arr["asmc"][i] = np.zeros((), dtype=uo.ASMC_DTYPE)
    env_arr["asmc"]["eta"][i] = g["reset_position"][k]
    env_arr["asmc"]["nu"][i] = g["reset_velocity"][k]
    env_arr["path_start"][i] = g["reset_path_start"][k]
    env_arr["path_end"][i] = g["reset_path_end"][k]
    env_arr["target"][i] = g["reset_target"][k]
    env_arr["max_action"][i] = g["reset_max_action"][k]
    env_arr["v_ref"][i] = g["reset_v_ref"][k]
    env_arr["n_obs"][i] = g["reset_n_obs"][k]
    env_arr["obs_x"][i] = g["reset_ox"][k]
    env_arr["obs_y"][i] = g["reset_oy"][k]
    env_arr["obs_r"][i] = g["reset_rad"][k]
    env_arr["progress"][i] = 0
    env_arr["steps"][i] = 0
    env_arr["ep_return"][i] = 0


def inject_gpu(state, i, g, k, keep_rows=()):
    """Put golden reset record k into column i of a UsvGpuVecEnv.get_state() dict (in place); keep_rows of
    USV_FIELD_DYN are left as they are ('usv-simple': last_action survives a reset)."""
    dyn, epi = state["dyn"], state["epi"]
    kept = [dyn[r, i] for r in keep_rows]
    dyn[:, i] = 0
    for r, v in zip(keep_rows, kept):
        dyn[r, i] = v
    dyn[0:3, i] = g["reset_position"][k]
    dyn[3:6, i] = g["reset_velocity"][k]
    epi[0:2, i] = g["reset_path_start"][k]
    epi[2:4, i] = g["reset_path_end"][k]
    epi[4, i] = g["reset_max_action"][k][0]
    epi[5, i] = g["reset_max_action"][k][2]
    epi[6, i] = g["reset_v_ref"][k]
    state["n_obs"][i] = g["reset_n_obs"][k]
    state["obstacles"][i, 0] = g["reset_ox"][k]
    state["obstacles"][i, 1] = g["reset_oy"][k]
    state["obstacles"][i, 2] = g["reset_rad"][k]
    state["steps"][i] = 0
    state["ep_return"][i] = 0


def oracle_to_gpu_state(envs, obstacles_dtype=np.float64):
    """Oracle env array -> UsvGpuVecEnv.set_state() dict."""
    n = len(envs)
    a = envs["asmc"]
    dyn = np.zeros((23, n))
    dyn[0:3] = a["eta"].T
    dyn[3:6] = a["nu"].T
    dyn[6:15] = a["last"].T
    dyn[15:19] = a["so"].T
    dyn[19:22] = a["aux"].T
    dyn[22] = envs["progress"]
    epi = np.zeros((7, n))
    epi[0:2] = envs["path_start"].T
    epi[2:4] = envs["path_end"].T
    epi[4] = envs["max_action"][:, 0]
    epi[5] = envs["max_action"][:, 2]
    epi[6] = envs["v_ref"]
    obst = np.zeros((n, 3, 32), dtype=obstacles_dtype)  # per-env rows: x[32], y[32], radius[32]
    obst[:, 0] = envs["obs_x"]
    obst[:, 1] = envs["obs_y"]
    obst[:, 2] = envs["obs_r"]
    return dict(dyn=dyn, epi=epi, obstacles=obst, n_obs=envs["n_obs"].astype(np.int32),
                steps=envs["steps"].astype(np.int32), episode=envs["episode"].astype(np.int32),
                ep_return=envs["ep_return"].copy())
