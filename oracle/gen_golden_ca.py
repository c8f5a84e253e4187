"""Golden rollouts of UsvAsmcCaEnv ('usv-asmc-ca-v0', gym_usv/envs/usv_asmc_ca_env.py) for the CA env kind.

TEST INFRASTRUCTURE ONLY.  Two things stand between the reference's file and a run, both handled here WITHOUT touching
the file:
  * usv_libs_py (compiled controller / model library, un-vendored, un-pinned) is absent.  SURVEY.md 8f rank 3 asks for
    the env "with the in-repo ASMC substituted"; the reference carries that substitution itself as a commented-out line
    (usv_asmc_ca_env.py:210).  The shim below gives usv_libs_py's call shape (:127-128, 196-205, 333, 385) to
    gym_usv.control.UsvAsmc: one update_controller_and_model_n(model, asmc, setpoint, 10) == one UsvAsmc.compute
    (10 sub-steps of 0.01 s) with action = [setpoint.velocity, setpoint.heading].
  * at HEAD the env's own step() does not type under numba (reset() reshapes obs_r to (-1, 1), :388-390, and the
    @njit _compute_sensor_distances is then asked for min(array, float)); NUMBA_DISABLE_JIT=1 runs the same function
    as plain Python, which is what the golden vectors record.
python oracle/gen_golden_ca.py  ->  tests/golden/ca_env.npz
"""
import os
import sys
import types

os.environ["NUMBA_DISABLE_JIT"] = "1"
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import numpy as np  # noqa: E402

import ref_harness as rh  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


class ShimDynamicModel:
    def __init__(self, x=0.0, y=0.0, psi=0.0):
        self.position = np.array([x, y, psi], dtype=np.float64)
        self.velocity = np.zeros(3)
        self.init_pose = self.position.copy()  # what reset() drew (:331-336), before its internal step([-1, 0])


class ShimASMC:
    def __init__(self, params=None):
        from gym_usv.control.usv_asmc import UsvAsmc
        self.impl = UsvAsmc()

    @staticmethod
    def defaultParams():
        return None


class ShimSetpoint:
    velocity = 0.0
    heading = 0.0


def shim_update(model, asmc, setpoint, n):
    assert n == 10
    pos, vel, infos = asmc.impl.compute([setpoint.velocity, setpoint.heading], model.position, model.velocity, False)
    model.position, model.velocity = np.array(pos, dtype=np.float64), np.array(vel, dtype=np.float64)
    out = types.SimpleNamespace(pose_x=pos[0], pose_y=pos[1], pose_psi=pos[2], vel_x=vel[0], vel_y=vel[1], vel_r=vel[2])
    return [out], [infos]


def load_ca_env():
    rh.install_stubs()
    lib = sys.modules["usv_libs_py"]
    lib.controller.ASMC, lib.controller.ASMCSetpoint = ShimASMC, ShimSetpoint
    lib.model.DynamicModel = ShimDynamicModel
    lib.utils.update_controller_and_model_n = shim_update
    from gym_usv.envs.usv_asmc_ca_env import UsvAsmcCaEnv
    return UsvAsmcCaEnv


def snapshot(env):
    n = int(env.num_obs)
    ox, oy, orad = (np.zeros(9) for _ in range(3))
    ox[:n], oy[:n], orad[:n] = env.obs_x.reshape(-1), env.obs_y.reshape(-1), env.obs_r.reshape(-1)
    m = env.asmc.impl
    return dict(position=np.array(env.position, np.float64), velocity=np.array(env.velocity, np.float64),
                target=np.array(env.target_point[:2], np.float64), n_obs=n, ox=ox, oy=oy, orad=orad,
                last=np.array(m.last, np.float64), so_filter=np.array(m.so_filter, np.float64),
                aux=np.array(m.aux_vars, np.float64), prev_action=np.array(env.action_history[-1], np.float64))


def rollout(Env, np_seed, act_seed, steps, hold, steer=False):
    np.random.seed(np_seed)
    rng = np.random.default_rng(act_seed)
    env = Env()  # __init__ ends with reset(), which ends with step([-1, 0])
    resets = [(0, snapshot(env), np.array(env.state, np.float32))]
    obs, rew = np.zeros((steps, 23), np.float32), np.zeros(steps)
    done, trunc = np.zeros(steps, np.uint8), np.zeros(steps, np.uint8)
    pos, vel, acts = np.zeros((steps, 3)), np.zeros((steps, 3)), np.zeros((steps, 2), np.float32)
    a = None
    for t in range(steps):
        if t % hold == 0:
            a = rng.uniform(-1, 1, 2).astype(np.float32)
            a[0] = abs(a[0])  # mostly forward, so that episodes reach targets / obstacles
            a[1] *= 0.15
        if steer:  # head for the target: action[1] is a heading offset relative to the course (usv_asmc.py:76)
            a = np.array([0.8, resets[-1][2][4] if t == resets[-1][0] else obs[t - 1][4]], np.float32)
        acts[t] = a
        o, r, d, tr, _ = env.step(a)
        obs[t], rew[t], done[t], trunc[t], pos[t], vel[t] = o, r, d, tr, env.position, env.velocity
        if d or tr:
            o, _ = env.reset()
            resets.append((t + 1, snapshot(env), np.array(o, np.float32)))
    out = dict(actions=acts, obs=obs, reward=rew, done=done, trunc=trunc, pos=pos, vel=vel,
               reset_step=np.array([r[0] for r in resets]), reset_obs=np.array([r[2] for r in resets]))
    for k in resets[0][1]:
        out["reset_" + k] = np.array([r[1][k] for r in resets])
    return out, len(resets), int(done.sum()), int(trunc.sum())


# reset(options=...) (usv_asmc_ca_env.py:358-372).  O0 is experiment_1_options() of tools/test_env.py:8-16 (that file
# imports pygame and cannot be imported here; its second layout has 78 circles, more than the 9 a drawn scene can hold).
# O1: circles inside the deletion radius of start_position / target_point (:374-399) and a start_position far from the
# drawn pose -- the boat does NOT start there (the model keeps the drawn pose, :336 vs :370-372).  O2: obstacles only;
# the deletion rule then runs against the drawn pose and the drawn target.
OPTION_CASES = [
    dict(obs_x=np.array([-6, 0, 6, 3, -3]), obs_y=np.array([0, 0, 0, 4, 4]), obs_r=np.array([1.5, 1.5, 1.5, 1.5, 1.5]),
         start_position=np.array([0, -8, 0]), target_point=np.array([0, 8, 0]), renderplots=False),
    dict(obs_x=np.array([2.0, 2.5, 14.0, 20.0, 21.0, 8.0]), obs_y=np.array([-7.0, -3.0, 2.0, 6.5, 3.0, -1.0]),
         obs_r=np.array([1.0, 1.2, 1.8, 1.0, 1.5, 2.0]), start_position=np.array([2.0, -8.5, 0.3]),
         target_point=np.array([20.0, 7.0, 0.0]), renderplots=False),
    dict(obs_x=np.array([0.0, 5.0, 10.0, 15.0, 20.0, 25.0, -5.0, 12.0, 3.0]),
         obs_y=np.array([-8.0, -7.0, 0.0, 4.0, 7.0, -7.5, -9.0, 8.0, 6.5]),
         obs_r=np.array([1.0, 1.5, 2.0, 1.0, 1.5, 2.0, 1.0, 1.5, 2.0]), renderplots=False),
]


def options_rollout(Env, np_seed, act_seed, steps, options):
    np.random.seed(np_seed)
    rng = np.random.default_rng(act_seed)
    env = Env()
    o, _ = env.reset(options=options)
    out = {"reset_" + k: np.asarray(v)[None] for k, v in snapshot(env).items()}  # same layout as rollout(): [resets, ...]
    out["reset_obs"] = np.array(o, np.float32)[None]
    out["reset_init_pose"] = env.model.init_pose.copy()[None]
    out["reset_step"] = np.array([0])
    obs, rew, done, trunc, pos, vel, acts = [], [], [], [], [], [], []
    for t in range(steps):
        a = np.array([0.8, (obs[-1] if obs else o)[4]], np.float32) if t % 3 else rng.uniform(-1, 1, 2).astype(np.float32)
        o2, r, d, tr, _ = env.step(a)
        acts.append(a), obs.append(np.array(o2, np.float32)), rew.append(r), done.append(d), trunc.append(tr)
        pos.append(np.array(env.position)), vel.append(np.array(env.velocity))
        if d or tr:
            break
    out.update(actions=np.array(acts, np.float32), obs=np.array(obs, np.float32), reward=np.array(rew),
               done=np.array(done, np.uint8), trunc=np.array(trunc, np.uint8), pos=np.array(pos), vel=np.array(vel))
    for k in ("obs_x", "obs_y", "obs_r", "start_position", "target_point"):
        if k in options:
            out["opt_" + k] = np.asarray(options[k], dtype=np.float64)
    return out


if __name__ == "__main__":
    Env = load_ca_env()
    out = {}
    for k, opt in enumerate(OPTION_CASES):
        o = options_rollout(Env, 60 + k, 70 + k, 400, opt)
        print(f"ca O{k}: kept {int(o['reset_n_obs'][0])} of {len(opt['obs_x'])} obstacles, {len(o['obs'])} steps, "
              f"done {int(o['done'].sum())} truncated {int(o['trunc'].sum())}")
        out.update({f"O{k}_{name}": v for name, v in o.items()})
    for k, (np_seed, act_seed, steps, hold, steer) in enumerate([(40, 41, 1500, 1, False), (42, 43, 3000, 40, False),
                                                                 (44, 45, 3000, 1, True)]):
        o, nres, nd, nt = rollout(Env, np_seed, act_seed, steps, hold, steer)
        print(f"ca R{k}: resets {nres} done {nd} truncated {nt}")
        out.update({f"R{k}_{name}": v for name, v in o.items()})
    np.savez_compressed(os.path.join(OUT, "ca_env.npz"), **out)
