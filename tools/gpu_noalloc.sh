#!/bin/bash
# usvgpu_sensors_noalloc: its test alone in a fresh process, then the whole GPU suite (which ends with it), then a timing
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q -s > gpurun_out/noalloc_alone.log 2>&1; echo "alone exit $?" | tee -a gpurun_out/noalloc_alone.log; tail -5 gpurun_out/noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/noalloc_suite.log 2>&1; echo "suite exit $?" | tee -a gpurun_out/noalloc_suite.log; tail -4 gpurun_out/noalloc_suite.log
timeout 300 python - <<'PY'
import ctypes as C, torch, numpy as np
import gym_usv_b200._native as nat
L = nat.lib(); dev = torch.device("cuda", 0); m = 8192
rng = np.random.default_rng(0)
pose = torch.tensor(np.stack([rng.uniform(0, 20, m), rng.uniform(0, 20, m), rng.uniform(-3, 3, m)]), device=dev)
obst = torch.tensor(np.concatenate([rng.uniform(0, 20, (32, 2, m)), rng.uniform(0.15, 0.5, (32, 1, m))], axis=1), device=dev).contiguous()
n_obs = torch.tensor(rng.integers(15, 30, m).astype(np.int32), device=dev)
out = torch.empty((128, m), dtype=torch.float64, device=dev)
p = lambda t: C.c_void_p(t.data_ptr())
for name in ("usvgpu_sensors", "usvgpu_sensors_noalloc"):
    f = getattr(L, name); st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    for _ in range(10): f(0, m, p(pose), p(obst), p(n_obs), p(out), st)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(200): f(0, m, p(pose), p(obst), p(n_obs), p(out), st)
    e1.record(); torch.cuda.synchronize()
    print(name, "%.2f us per call of %d envs" % (e0.elapsed_time(e1) / 200 * 1e3, m))
PY
