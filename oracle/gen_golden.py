"""Generate tests/golden/ (npz files) by running the UNMODIFIED reference in this container.

TEST INFRASTRUCTURE ONLY.  Run once here (the reference is not on the GPU box):

    python oracle/gen_golden.py

The reference publishes no golden vectors; these files are outputs
of the reference's own code (via oracle/ref_harness.py) and are what pins oracle/usv_oracle.c.
Configs follow SURVEY.md section 8d (1a-1d).
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_harness as rh  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")
MAX_OBS = 32


def action_stream(n, seed=1):
    """config 1a/1b: default_rng(1).uniform([0.2,-1],[1,1]) cast to f32 (the Box dtype, simple_env.py:30)."""
    return np.random.default_rng(seed).uniform([0.2, -1.0], [1.0, 1.0], size=(n, 2)).astype(np.float32)


# -----------------------------------------------------------------------------------------
def gen_asmc(R):
    out = {}
    fixed = np.array([[0.0, 0.0], [10.0, 0.0], [0.0, 10.0]])  # tests/test_usv_asmc.py:13,23,35
    checkpoints = np.array([1, 3, 10, 100, 500, 1000])
    fpos = np.zeros((3, len(checkpoints), 3))
    fvel = np.zeros_like(fpos)
    for k, act in enumerate(fixed):
        asmc = R.UsvAsmc()
        p, v = np.zeros(3), np.zeros(3)
        for c in range(1, 1001):
            p, v, _ = asmc.compute(act, p, v, False)
            if c in checkpoints:
                j = int(np.where(checkpoints == c)[0][0])
                fpos[k, j], fvel[k, j] = p, v
    out.update(fixed_actions=fixed, fixed_checkpoints=checkpoints, fixed_pos=fpos, fixed_vel=fvel)

    # anchor from SURVEY section 8a: zero state, action [1.0, 0.5]
    asmc = R.UsvAsmc()
    p, v = np.zeros(3), np.zeros(3)
    anchor = []
    for c in range(3):
        p, v, infos = asmc.compute(np.array([1.0, 0.5]), p, v, False)
        anchor.append(np.hstack([p, v, infos[-1]["tport"], infos[-1]["tstbd"]]))
    out["anchor_action"] = np.array([1.0, 0.5])
    out["anchor"] = np.array(anchor)

    for name, perturb, n in (("rand", False, 1000), ("perturb", True, 300)):
        acts = action_stream(n)
        asmc = R.UsvAsmc()
        p, v = np.zeros(3), np.zeros(3)
        pos, vel, thr = np.zeros((n, 3)), np.zeros((n, 3)), np.zeros((n, 4))
        for c in range(n):
            p, v, infos = asmc.compute(acts[c], p, v, perturb)
            pos[c], vel[c] = p, v
            thr[c] = [infos[-1]["psi_d"], infos[-1]["u_d"], infos[-1]["tport"], infos[-1]["tstbd"]]
        out.update({f"{name}_actions": acts, f"{name}_pos": pos, f"{name}_vel": vel, f"{name}_info": thr,
                    f"{name}_last": np.asarray(asmc.last, dtype=np.float64),
                    f"{name}_so_filter": np.asarray(asmc.so_filter, dtype=np.float64),
                    f"{name}_aux": np.asarray(asmc.aux_vars, dtype=np.float64)})
    np.savez_compressed(os.path.join(OUT, "asmc_compute.npz"), **out)
    print("asmc_compute:", {k: v.shape for k, v in out.items()})


# -----------------------------------------------------------------------------------------
def gen_sensors(R):
    rng = np.random.default_rng(7)
    cases = []
    for S in (128, 16):
        pitch = ((2 / 3) * (2 * np.pi)) / S
        for c in range(96 if S == 128 else 16):
            n = int(rng.integers(1, 30))
            pos = np.array([rng.uniform(0, 20), rng.uniform(0, 20), rng.uniform(-7, 7)])
            ox, oy = rng.uniform(0, 20, n), rng.uniform(0, 20, n)
            rad = rng.uniform(0.15, 0.5, n)
            if c % 8 == 1:  # boat inside an obstacle (terminal-step case, negative readings)
                ox[0], oy[0] = pos[0] + 0.1 * rad[0], pos[1] - 0.2 * rad[0]
            if c % 8 == 2:  # overlapping obstacles in a row (min-key rule)
                d = np.array([np.cos(pos[2]), np.sin(pos[2])])
                for j in range(min(n, 4)):
                    ox[j], oy[j] = pos[:2] + d * (1.0 + 0.4 * j) + rng.normal(0, 0.15, 2)
                    rad[j] = rng.uniform(0.3, 0.5)
            if c % 8 == 3:  # very close obstacle directly behind (seen at both fan edges)
                d = np.array([np.cos(pos[2] + np.pi), np.sin(pos[2] + np.pi)])
                ox[0], oy[0] = pos[:2] + d * 0.52
                rad[0] = 0.5
            dist = np.hypot(ox - pos[0], oy - pos[1]) - rad
            sens = R.UsvAsmcCaEnv.compute_sensor_measurments(pos, S, 100, rad, pitch, ox.reshape(-1, 1),
                                                             oy.reshape(-1, 1), n, dist)
            cases.append((S, pos, n, ox, oy, rad, sens[:, 1].copy()))
    m = len(cases)
    out = dict(S=np.array([c[0] for c in cases]), pos=np.array([c[1] for c in cases]),
               n=np.array([c[2] for c in cases]), ox=np.zeros((m, MAX_OBS)), oy=np.zeros((m, MAX_OBS)),
               rad=np.zeros((m, MAX_OBS)), dist=np.full((m, 128), np.nan))
    for i, c in enumerate(cases):
        out["ox"][i, :c[2]], out["oy"][i, :c[2]], out["rad"][i, :c[2]] = c[3], c[4], c[5]
        out["dist"][i, :c[0]] = c[6]
    np.savez_compressed(os.path.join(OUT, "sensors.npz"), **out)
    print("sensors:", m, "cases; hits:", int(np.sum(out["dist"] < 100)))


# -----------------------------------------------------------------------------------------
def _reset_record(env, obs):
    n = int(env.obstacle_n)
    ox, oy, rad = np.zeros(MAX_OBS), np.zeros(MAX_OBS), np.zeros(MAX_OBS)
    ox[:n], oy[:n] = env.obstacle_positions[:, 0], env.obstacle_positions[:, 1]
    rad[:n] = env.obstacle_radius
    return dict(position=np.array(env.position, dtype=np.float64), velocity=np.array(env.velocity, dtype=np.float64),
                path_start=np.array(env.path_start), path_end=np.array(env.path_end),
                target=np.array(env.target_position), max_action=np.array(env.max_action),
                v_ref=float(env.reference_velocity), n_obs=n, ox=ox, oy=oy, rad=rad, obs=np.array(obs))


def gen_simple_env(R, name, seeds, steps, ignore_obstacles, cls="UsvSimpleASMCEnv", limit=1000):
    acts = action_stream(steps)
    E = len(seeds)
    obs = np.zeros((E, steps, 143), np.float32)
    rew = np.zeros((E, steps))
    term = np.zeros((E, steps), np.uint8)
    trunc = np.zeros((E, steps), np.uint8)
    pos = np.zeros((E, steps, 3))
    vel = np.zeros((E, steps, 3))
    prog = np.zeros((E, steps))
    resets = []  # (env, step index BEFORE which the reset happened, record)
    for e, seed in enumerate(seeds):
        env = getattr(R, cls)()
        env.ignore_obstacles = ignore_obstacles
        tl = rh.TimeLimit(env, limit)  # gym_usv/__init__.py:24-34
        o, _ = tl.reset(seed=seed)
        resets.append((e, 0, _reset_record(env, o)))
        for t in range(steps):
            o, r, te, tr, _ = tl.step(acts[t])
            obs[e, t], rew[e, t], term[e, t], trunc[e, t] = o, r, te, tr
            pos[e, t], vel[e, t], prog[e, t] = env.position, env.velocity, env.progress
            if te or tr:
                o, _ = tl.reset()  # seed=None: the env's generator continues (SB3 auto-reset)
                resets.append((e, t + 1, _reset_record(env, o)))
    out = dict(seeds=np.array(seeds), actions=acts, obs=obs, reward=rew, terminated=term, truncated=trunc, pos=pos,
               vel=vel, progress=prog, reset_env=np.array([r[0] for r in resets]),
               reset_step=np.array([r[1] for r in resets]))
    for k in resets[0][2]:
        out["reset_" + k] = np.array([r[2][k] for r in resets])
    np.savez_compressed(os.path.join(OUT, name), **out)
    print(name, "envs", E, "resets", len(resets), "terminated", int(term.sum()), "truncated", int(trunc.sum()))


def gen_reset_samples(R, n=1500):
    env = R.UsvSimpleASMCEnv()
    keys = ("position", "velocity", "path_start", "path_end", "target", "max_action", "v_ref", "n_obs")
    rec = {k: [] for k in keys}
    rec["rad_mean"], rec["obs_mean"], rec["obs0"] = [], [], []
    for i in range(n):
        o, _ = env.reset(seed=10_000 + i)
        r = _reset_record(env, o)
        for k in keys:
            rec[k].append(r[k])
        rec["rad_mean"].append(r["rad"][:r["n_obs"]].mean())
        rec["obs_mean"].append([r["ox"][:r["n_obs"]].mean(), r["oy"][:r["n_obs"]].mean()])
        rec["obs0"].append(np.array(o[:15]))
    np.savez_compressed(os.path.join(OUT, "reset_samples.npz"), **{k: np.array(v) for k, v in rec.items()})
    print("reset_samples:", n)


# -----------------------------------------------------------------------------------------
def _legacy_rollout(R, np_seed, act_seed, steps, full, hold=1, cls="UsvAsmcEnv"):
    """UsvAsmcEnv (usv-asmc-v0) or one of its float64 siblings: np.random.seed(np_seed), scalar f32 actions U(+-pi/2)
    each held for `hold` steps, auto-reset on done."""
    np.random.seed(np_seed)
    acts = np.random.default_rng(act_seed).uniform(-np.pi / 2, np.pi / 2, (steps + hold - 1) // hold)
    acts = np.repeat(acts, hold)[:steps].astype(np.float32)
    env = getattr(R, cls)()
    f64 = cls != "UsvAsmcEnv"  # the siblings return and keep float64 everywhere
    resets = []

    def rec():
        return dict(state=np.array(env.state, np.float64), velocity=np.array(env.velocity, np.float64),
                    position=np.array(env.position, np.float64), aux=np.array(env.aux_vars, np.float64),
                    last=np.array(env.last, np.float64), target=np.array(env.target, np.float64))

    o = env.reset()
    resets.append((0, rec(), o))
    na, nl = len(env.aux_vars), len(env.last)
    obs = np.zeros((steps, 6), np.float64 if f64 else np.float32)
    rew, done = np.zeros(steps), np.zeros(steps, np.uint8)
    pos, vel, aux, last = np.zeros((steps, 3)), np.zeros((steps, 3)), np.zeros((steps, na)), np.zeros((steps, nl))
    for t in range(steps):
        o, r, d, _ = env.step(acts[t])
        obs[t], rew[t], done[t] = o, r, d
        pos[t], vel[t], aux[t], last[t] = env.position, env.velocity, env.aux_vars, env.last
        if d:
            o = env.reset()
            resets.append((t + 1, rec(), o))
    out = dict(actions=acts, obs=obs, reward=rew, done=done, pos=pos, vel=vel,
               reset_step=np.array([r[0] for r in resets]), reset_obs=np.array([r[2] for r in resets]))
    if full:
        out.update(aux=aux, last=last)
    elif f64:
        pass
    else:  # after a step the arrays hold float32 values (usv_asmc_env.py:247-251): store them as such
        out["pos"], out["vel"] = pos.astype(np.float32), vel.astype(np.float32)
        assert np.array_equal(out["pos"].astype(np.float64), pos) and np.array_equal(out["vel"].astype(np.float64), vel)
    for k in resets[0][1]:
        out["reset_" + k] = np.array([r[1][k] for r in resets])
    return out, len(resets), int(done.sum())


def gen_legacy(R):
    """config 1d: 1000 steps from np.random.seed(0) with every array recorded, plus two longer rollouts (keys 'L0_*':
    2000 steps, a new action every step; 'L1_*': 6000 steps, actions held for 250 steps so that the vessel runs into
    the |x| > 30 / |ye| > 10 terminations and the float64 first step after each reset)."""
    out, nres, nd = _legacy_rollout(R, 0, 2, 1000, full=True)
    print("legacy: resets", nres, "done", nd)
    for k, (np_seed, act_seed, steps, hold) in enumerate([(1, 3, 2000, 1), (2, 4, 6000, 250)]):
        o, nres, nd = _legacy_rollout(R, np_seed, act_seed, steps, full=False, hold=hold)
        print(f"legacy L{k}: resets", nres, "done", nd)
        out.update({f"L{k}_{name}": v for name, v in o.items()})
    np.savez_compressed(os.path.join(OUT, "legacy_asmc_env.npz"), **out)


def gen_legacy_siblings(R):
    """'usv-asmc-ye-int-v0' and 'usv-pid-v0' (gym_usv/__init__.py:8-16): 1500 steps with a new action every step and
    every array recorded, plus 6000 steps with actions held for 250 / 500 steps (keys 'L1_*') so that the |ye| > 10 /
    x < -10 terminations and the resets after them are on file."""
    for name, cls, s0, l1 in (("ye_int", "UsvAsmcYeIntEnv", 10, (12, 13, 6000, 250)),
                              ("pid", "UsvPidEnv", 20, (24, 25, 6000, 500))):
        out, nres, nd = _legacy_rollout(R, s0, s0 + 1, 1500, full=True, cls=cls)
        print(f"legacy {name}: resets", nres, "done", nd)
        o, nres, nd = _legacy_rollout(R, l1[0], l1[1], l1[2], full=False, hold=l1[3], cls=cls)
        print(f"legacy {name} L1: resets", nres, "done", nd)
        out.update({f"L1_{k}": v for k, v in o.items()})
        np.savez_compressed(os.path.join(OUT, f"legacy_{name}_env.npz"), **out)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    R = rh.load_reference()
    which = sys.argv[1:] or ["asmc", "sensors", "simple", "simple_noobs", "resets", "legacy", "legacy_siblings", "kinematic"]
    if "asmc" in which:
        gen_asmc(R)
    if "sensors" in which:
        gen_sensors(R)
    if "simple" in which:
        gen_simple_env(R, "simple_asmc_env.npz", seeds=list(range(8)), steps=1000, ignore_obstacles=False)
    if "simple_noobs" in which:
        gen_simple_env(R, "simple_asmc_env_noobs.npz", seeds=[100, 101, 102, 103], steps=1000, ignore_obstacles=True)
    if "resets" in which:
        gen_reset_samples(R)
    if "legacy" in which:
        gen_legacy(R)
    if "legacy_siblings" in which:
        gen_legacy_siblings(R)
    if "kinematic" in which:  # 'usv-simple' = UsvSimpleEnv itself, TimeLimit 500 (gym_usv/__init__.py:24-28)
        gen_simple_env(R, "simple_kinematic_env.npz", seeds=[200, 201, 202], steps=700, ignore_obstacles=False,
                       cls="UsvSimpleEnv", limit=500)
