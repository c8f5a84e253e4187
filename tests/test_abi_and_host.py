"""CPU-only checks: the C-ABI library loads and exports every symbol include/usvgpu.h declares,
struct layouts agree, the host-side mirror (spaces, infos, sharding) behaves, and without a GPU
the product path fails loudly instead of falling back."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_functions():
    src = open(os.path.join(ROOT, "include", "usvgpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(usvgpu_[a-z_0-9]+)\s*\(", src)))


def test_library_builds_loads_and_exports_every_declared_symbol():
    from gym_usv_b200 import _native as nat
    path = nat.build()
    L = C.CDLL(path)
    names = _header_functions()
    assert len(names) >= 14
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/usvgpu.h but not exported"
    assert set(nat.EXPORTS) == set(names), set(nat.EXPORTS) ^ set(names)
    L.usvgpu_version.restype = C.c_int
    assert L.usvgpu_version() == 100


def test_struct_layouts_match_header():
    from gym_usv_b200 import _native as nat
    assert C.sizeof(nat.UsvConfig) == 8 * 4
    assert C.sizeof(nat.UsvStats) == 8 * 8
    hdr = open(os.path.join(ROOT, "include", "usvgpu.h")).read()
    assert f"#define USVGPU_OBS_DIM {nat.OBS_DIM}" in hdr
    assert f"#define USV_DYN_ROWS {nat.DYN_ROWS}" in hdr and f"#define USV_EPI_ROWS {nat.EPI_ROWS}" in hdr
    assert len(nat.DYN_NAMES) == nat.DYN_ROWS and len(nat.EPI_NAMES) == nat.EPI_ROWS


def test_no_cpu_fallback_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    import gym_usv_b200 as m
    with pytest.raises(m.UsvError):
        m.UsvGpuVecEnv(4)
    with pytest.raises(m.UsvError):
        m.asmc_compute(torch.zeros((23, 1), dtype=torch.float64), torch.zeros((1, 2), dtype=torch.float64))
    # straight through the C ABI: create must refuse, with a message, and never hand out a handle
    from gym_usv_b200 import _native as nat
    cfg = nat.UsvConfig(0, 0, 0, 0, 1000, 1, 0, 0)
    h = C.c_void_p()
    rc = nat.lib().usvgpu_create(C.byref(cfg), 0, 16, 0, 0, C.byref(h))
    assert rc == -3 and not h.value
    assert b"no CUDA device" in nat.lib().usvgpu_last_error(None)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "gym_usv_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                s = open(os.path.join(dp, f)).read()
                assert "usv_oracle" not in s and "ref_harness" not in s, f


def test_spaces_match_reference_definitions():
    from gym_usv_b200 import spaces
    o, a = spaces.simple_observation_space(), spaces.simple_action_space()
    assert o.shape == (143,) and o.dtype == np.float32 and float(o.low.min()) == -1 and float(o.high.max()) == 1
    np.testing.assert_array_equal(a.low, np.array([0.2, -1], np.float32))
    np.testing.assert_array_equal(a.high, np.array([1, 1], np.float32))
    s = a.sample()
    assert a.contains(s) and s.dtype == np.float32
    lo = spaces.legacy_observation_space()
    assert lo.shape == (6,) and spaces.legacy_action_space().shape == (1,)


def test_lazy_infos():
    from gym_usv_b200.vec_env import LazyInfos
    tob = np.arange(6, dtype=np.float32).reshape(2, 3)
    li = LazyInfos(5, [1, 4], np.array([True, False]), np.array([1.5, -2.0]), np.array([10.0, 3.0]), tob)
    a = {"TimeLimit.truncated": True, "episode": {"r": 1.5, "l": 10}}
    b = {"TimeLimit.truncated": False, "episode": {"r": -2.0, "l": 3}}
    assert len(li) == 5 and li[0] == {} and li[2] == {}
    for got, want, row in ((li[1], a, tob[0]), (li[4], b, tob[1]), (li[-1], b, tob[1])):
        assert np.array_equal(got.pop("terminal_observation"), row) and got == want
    assert li[0:1] == [{}] and [i for i, _ in li.done_items()] == [1, 4]
    assert "terminal_observation" not in LazyInfos(5, [1], [True], [0.0], [1.0])[1]  # nothing was reset: no SB3 key
    with pytest.raises(IndexError):
        li[5]


def test_shard_ranges_cover_and_are_contiguous():
    from gym_usv_b200.sharding import owner_of, shard_range
    for n, w in ((1048576, 8), (1048576, 4), (65536, 2), (10, 3), (7, 8), (0, 2)):
        hi_prev = 0
        for r in range(w):
            lo, hi = shard_range(n, r, w)
            assert lo == hi_prev and hi >= lo
            hi_prev = hi
            for e in {lo, hi - 1} if hi > lo else ():
                assert owner_of(e, n, w) == r
        assert hi_prev == n
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


_WORKER = r"""
import os, sys
sys.path.insert(0, {root!r})
import torch, torch.distributed as dist
from gym_usv_b200.sharding import shard_range, reduce_episode_stats, STAT_KEYS
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", rank=rank, world_size=world)
lo, hi = shard_range(1001, rank, world)
stats = dict.fromkeys(STAT_KEYS, 0)
stats.update(return_sum=1.5 * (rank + 1), length_sum=10.0 * (rank + 1), episodes=rank + 1, env_steps=hi - lo)
tot = reduce_episode_stats(stats)
assert tot["env_steps"] == 1001, tot
assert tot["episodes"] == sum(range(1, world + 1)), tot
assert abs(tot["return_sum"] - 1.5 * sum(range(1, world + 1))) < 1e-12
dist.barrier()
dist.destroy_process_group()
print("OK", rank)
"""


def test_stats_reduce_world_size_2_gloo(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(_WORKER.format(root=ROOT))
    procs = []
    port = 29500 + (os.getpid() % 2000)
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    for p in procs:
        out, _ = p.communicate(timeout=180)
        assert p.returncode == 0 and "OK" in out, out


def test_bench_reference_arm_prints_one_json_line():
    """bench.py's contract for the reference arm (the only arm that runs without a GPU): exactly one JSON line on
    stdout with the driver's keys, cpu_baseline describing the run and a zero-copy e2e block."""
    import json
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "3",
                        "--cpu-budget-s", "0.5", "--ref-envs", "64"], capture_output=True, text=True, timeout=300, cwd=root)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, r.stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "env_steps_per_sec" and d["unit"] == "env-steps/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["value"] > 0
    assert d["cpu_baseline"]["kind"] in ("port", "reference") and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"]
