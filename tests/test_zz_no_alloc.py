"""usvgpu_sensors_noalloc (include/usvgpu.h; no counterpart in the reference):
the sensor entry point whose scratch is a fixed static device array, called through the C ABI and compared with the
existing usvgpu_sensors on the same inputs -- two calls in a row (a large batch, then a smaller one that finds the first
call's rows still in the scratch), the size limit, and the oracle on top.  Collected last on purpose
(file name)."""
import ctypes as C

import numpy as np
import pytest
import torch

import usv_oracle as uo

pytestmark = pytest.mark.gpu

ATOL = 1e-9  # the tolerance of the sensor tests in tests/test_gpu_parity.py (float64 distances in metres)


@pytest.fixture(scope="module")
def nat():
    if not torch.cuda.is_available():
        pytest.fail("gpu-marked test collected without a CUDA device")
    import gym_usv_b200._native as n
    n.lib()
    return n


def _scenes(seed, m):
    """Random scenes in the [32][3][m] / [3][m] layout of the unit entry points: ragged obstacle counts (0..32), boats
    inside circles now and then, large |psi|, junk in the unused slots (must be ignored)."""
    rng = np.random.default_rng(seed)
    pose = np.stack([rng.uniform(0, 20, m), rng.uniform(0, 20, m), rng.uniform(-9, 9, m)])
    obst = np.empty((32, 3, m))
    obst[:, 0], obst[:, 1] = rng.uniform(0, 20, (32, m)), rng.uniform(0, 20, (32, m))
    obst[:, 2] = rng.uniform(0.15, 0.5, (32, m))
    inside = rng.random(m) < 0.02
    obst[0, 0, inside], obst[0, 1, inside], obst[0, 2, inside] = pose[0, inside] + 0.05, pose[1, inside], 0.4
    n_obs = rng.integers(0, 33, m).astype(np.int32)
    n_obs[inside] = np.maximum(n_obs[inside], 1)
    return pose, obst, n_obs


def _call(nat, name, pose, obst, n_obs, out):
    p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    st = C.c_void_p(torch.cuda.current_stream(pose.device).cuda_stream)
    return getattr(nat.lib(), name)(0, pose.shape[1], p(pose), p(obst), p(n_obs), p(out), st)


def _both(nat, seed, m):
    dev = torch.device("cuda", 0)
    pose_h, obst_h, n_h = _scenes(seed, m)
    pose, obst, n_obs = (torch.tensor(a, device=dev).contiguous() for a in (pose_h, obst_h, n_h))
    want = torch.full((nat.SENSORS, m), -1.0, dtype=torch.float64, device=dev)
    got = torch.full((nat.SENSORS, m), -2.0, dtype=torch.float64, device=dev)
    assert _call(nat, "usvgpu_sensors", pose, obst, n_obs, want) == 0
    assert _call(nat, "usvgpu_sensors_noalloc", pose, obst, n_obs, got) == 0
    torch.cuda.synchronize(dev)
    return got.cpu().numpy(), want.cpu().numpy(), (pose_h, obst_h, n_h)


def test_noalloc_sensors_match_the_allocating_entry_point_on_two_calls_in_a_row(nat):
    worst, hits = 0.0, 0
    # second call smaller than the first (stale rows beyond it and under it), third at the limit, fourth tiny and ragged
    for seed, m in ((1, 6000), (2, 777), (3, nat.NOALLOC_MAX_ENVS), (4, 5)):
        got, want, (pose, obst, n_obs) = _both(nat, seed, m)
        assert np.array_equal(got < 100, want < 100), (seed, m)
        np.testing.assert_allclose(got, want, rtol=0, atol=ATOL)
        worst = max(worst, float(np.abs(got - want).max()))
        hits += int((want < 100).sum())
        if m <= 777:  # and the checker on top, on the small batches
            for i in range(0, m, 97):
                k = int(n_obs[i])
                if k == 0:
                    assert np.all(got[:, i] == 100.0)
                    continue
                ref = uo.sensors(pose[:, i], obst[:k, 2, i], obst[:k, 0, i], obst[:k, 1, i], sensor_count=128)
                np.testing.assert_allclose(got[:, i], ref, rtol=0, atol=ATOL)
    print(f"no-alloc sensors vs usvgpu_sensors: worst |diff| {worst:.3g} m over {hits} hits")
    assert hits > 10000


def test_noalloc_sensors_refuse_a_batch_beyond_the_fixed_scratch(nat):
    dev = torch.device("cuda", 0)
    m = nat.NOALLOC_MAX_ENVS + 1
    pose_h, obst_h, n_h = _scenes(5, m)
    pose, obst, n_obs = (torch.tensor(a, device=dev).contiguous() for a in (pose_h, obst_h, n_h))
    out = torch.full((nat.SENSORS, m), -3.0, dtype=torch.float64, device=dev)
    rc = _call(nat, "usvgpu_sensors_noalloc", pose, obst, n_obs, out)
    assert rc != 0
    msg = nat.lib().usvgpu_last_error(None).decode()
    assert str(nat.NOALLOC_MAX_ENVS) in msg and "exceeds" in msg, msg
    torch.cuda.synchronize(dev)
    assert bool((out == -3.0).all())  # nothing was launched
    # and the entry point still works afterwards
    got, want, _ = _both(nat, 6, 300)
    np.testing.assert_allclose(got, want, rtol=0, atol=ATOL)
