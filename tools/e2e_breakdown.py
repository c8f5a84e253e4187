#!/usr/bin/env python
"""Where the time of one host-face step goes (65 536 envs): raw D2H bandwidth, flags latency, whole step."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gym_usv_b200 as usv

n = 65536
dev = torch.device("cuda", 0)
src = torch.zeros((n, 143), device=dev)
dst = torch.zeros((n, 143)).pin_memory()
for _ in range(3):
    dst.copy_(src, non_blocking=True)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(20):
    dst.copy_(src, non_blocking=True)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 20
print(f"raw D2H {src.numel()*4/1e6:.1f} MB pinned: {dt*1e6:.0f} us = {src.numel()*4/dt/1e9:.1f} GB/s")

for mapped, chunks in ((False, 1), (False, 2), (False, 3), (False, 4), (True, 1)):
    env = usv.UsvGpuVecEnv(n, device=0, seed=0, obstacles_f32=True, info_mode="lazy", mapped_obs=mapped,
                           host_chunks=chunks)
    env.reset()
    a = np.random.default_rng(0).uniform([0.2, -1], [1, 1], (n, 2)).astype(np.float32)
    for _ in range(50):
        env.step(a)
    T = {"async": 0.0, "flags": 0.0, "wait": 0.0}
    K = 50
    for _ in range(K):
        t0 = time.perf_counter()
        env.step_async(a)
        t1 = time.perf_counter()
        env._lib.usvgpu_flags_sync(env._h)
        t2 = time.perf_counter()
        env.step_wait()
        t3 = time.perf_counter()
        T["async"] += t1 - t0; T["flags"] += t2 - t1; T["wait"] += t3 - t2
    print("mapped" if mapped else f"staged x{chunks}", {k: f"{v/K*1e6:.0f} us" for k, v in T.items()},
          f"total {sum(T.values())/K*1e6:.0f} us -> {n*K/sum(T.values())/1e6:.1f} M env-steps/s")
    env.close()
