"""The C oracle against the LIVE reference (oracle/ref_harness.py imports the unmodified sources from /root/reference).

The golden files under tests/golden/ are outputs of the same reference at the time they were generated; this test keeps
the oracle, the goldens and the reference from drifting apart silently: fresh seeds and actions every run of the suite in
the build container, skipped where /root/reference does not exist (the GPU box).  CPU only.
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import ref_harness as rh  # noqa: E402
import usv_oracle as uo  # noqa: E402

pytestmark = pytest.mark.skipif(not rh.reference_available(), reason="/root/reference is not present on this machine")


@pytest.fixture(scope="module")
def R():
    return rh.load_reference()


def test_asmc_compute_matches_the_live_reference(R):
    """gym_usv/control/usv_asmc.py:53-244, 150 calls of fresh random actions, with and without the disturbance."""
    rng = np.random.default_rng(2024)
    for perturb in (False, True):
        ref, o = R.UsvAsmc(), uo.OracleAsmc(1)
        pos, vel = np.zeros(3), np.zeros(3)
        for c in range(150):
            a = rng.uniform([0.2, -1.0], [1.0, 1.0]).astype(np.float32).astype(np.float64)
            pos, vel, infos = ref.compute(a, pos, vel, perturb)
            info = o.compute_one(0, a, do_perturb=perturb)
            np.testing.assert_allclose(o.s["eta"][0], pos, rtol=1e-8, atol=1e-10, err_msg=f"call {c}")
            np.testing.assert_allclose(o.s["nu"][0], vel, rtol=1e-8, atol=1e-10)
            last = infos[-1]
            np.testing.assert_allclose(info[2:], [last["tport"], last["tstbd"]], rtol=1e-6, atol=1e-7)


def test_sensor_model_matches_the_live_reference(R):
    """UsvAsmcCaEnv.compute_sensor_measurments (usv_asmc_ca_env.py:411-461, 500-519) on fresh random scenes."""
    rng = np.random.default_rng(7)
    S, pitch = 128, ((2 / 3) * (2 * np.pi)) / 128
    hits = 0
    for case in range(60):
        n = int(rng.integers(1, 30))
        pos = np.array([rng.uniform(0, 20), rng.uniform(0, 20), rng.uniform(-7, 7)])
        ox, oy, rad = rng.uniform(0, 20, n), rng.uniform(0, 20, n), rng.uniform(0.15, 0.5, n)
        dist = np.hypot(ox - pos[0], oy - pos[1]) - rad
        want = R.UsvAsmcCaEnv.compute_sensor_measurments(pos, S, 100, rad, pitch, ox.reshape(-1, 1), oy.reshape(-1, 1), n,
                                                         dist)[:, 1]
        got = uo.sensors(pos, rad, ox, oy, sensor_count=S)
        assert np.array_equal(got < 100, want < 100), case
        np.testing.assert_allclose(got, want, rtol=0, atol=1e-9)
        hits += int(np.sum(want < 100))
    assert hits > 200


@pytest.mark.parametrize("cls,kinematic", [("UsvSimpleASMCEnv", False), ("UsvSimpleEnv", True)])
def test_simple_env_step_matches_the_live_reference(R, cls, kinematic):
    """UsvSimpleASMCEnv / UsvSimpleEnv .step (simple_env_asmc.py:18-27, simple_env.py:310-346): the reference's reset
    state installed in the oracle (PCG64 is not replicated), then lock-step until the first done."""
    rng = np.random.default_rng(99)
    for seed in (11, 12):
        env = getattr(R, cls)()
        o, _ = env.reset(seed=seed)
        ov = uo.OracleVecEnv(1, auto_reset=False, kinematic=kinematic, max_episode_steps=0)
        e = ov.envs
        e["asmc"][0] = np.zeros((), dtype=uo.ASMC_DTYPE)
        e["asmc"]["eta"][0], e["asmc"]["nu"][0] = env.position, env.velocity
        e["path_start"][0], e["path_end"][0], e["target"][0] = env.path_start, env.path_end, env.target_position
        e["max_action"][0], e["v_ref"][0] = env.max_action, env.reference_velocity
        n = int(env.obstacle_n)
        e["n_obs"][0] = n
        e["obs_x"][0, :n], e["obs_y"][0, :n] = env.obstacle_positions[:, 0], env.obstacle_positions[:, 1]
        e["obs_r"][0, :n] = env.obstacle_radius
        ob = np.zeros(uo.OBS_DIM, np.float32)
        uo.lib().usv_oracle_simple_reset_obs(e[0:1].ctypes.data, ob.ctypes.data)
        np.testing.assert_allclose(ob, o, rtol=0, atol=1e-6)
        for t in range(200):
            a = rng.uniform([0.2, -1.0], [1.0, 1.0]).astype(np.float32)
            ro, rr, rterm, rtrunc, _ = env.step(a)
            oo, orew, oterm, otrunc = ov.step(a[None])
            assert bool(oterm[0]) == bool(rterm) and bool(otrunc[0]) == bool(rtrunc), (seed, t)
            np.testing.assert_allclose(oo[0], ro, rtol=0, atol=1e-6, err_msg=f"seed {seed} step {t}")
            np.testing.assert_allclose(orew[0], rr, rtol=0, atol=1e-9)
            np.testing.assert_allclose(e["asmc"]["eta"][0], env.position, rtol=0, atol=1e-9)
            if rterm or rtrunc:
                break
