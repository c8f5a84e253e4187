"""A/B of the overlapped step (usvgpu_set_overlap) against the plain two-launch step on one GPU.
python tools/time_overlap.py N_ENVS "chunk,sens,dyn" ["chunk,sens,dyn" ...]      ("0" = plain step)
Three rotating shards (working set > L2), preroll to the steady-state episode mix, K steps captured once into a CUDA graph
per configuration, median of R replays timed with CUDA events; configurations are interleaved twice."""
import os
import statistics
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import gym_usv_b200 as usv  # noqa: E402

n = int(sys.argv[1])
configs = [tuple(int(x) for x in c.split(",")) for c in sys.argv[2:]] or [(0,)]
K = max(6, min(60, (1 << 22) // n * 3))
R = 7
dev = torch.device("cuda", 0)
envs = [usv.UsvGpuVecEnv(n, device=0, seed=s, obstacles_f32=True, info_mode="lazy") for s in range(3)]
gen = torch.Generator(device=dev).manual_seed(0)
acts = (torch.rand((8, n, 2), generator=gen, device=dev) * torch.tensor([0.8, 2.0], device=dev) +
        torch.tensor([0.2, -1.0], device=dev)).contiguous()
for e in envs:
    e.reset_tensor()
for k in range(3 * 300):  # preroll
    envs[k % 3].step_tensor(acts[k % 8])
torch.cuda.synchronize(dev)


def measure(cfg):
    for e in envs:
        e.set_overlap(*cfg) if cfg[0] else e.set_overlap(0)
    for k in range(3):
        envs[k].step_tensor(acts[k])
    torch.cuda.synchronize(dev)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, capture_error_mode="thread_local"):
        for k in range(K):
            envs[k % 3].step_tensor(acts[k % 8])
    g.replay()
    torch.cuda.synchronize(dev)
    ms = []
    for _ in range(R):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize(dev)
        ms.append(e0.elapsed_time(e1))
    return statistics.median(ms) / K * 1e3


for rep in range(2):
    for cfg in configs:
        us = measure(cfg)
        print(f"n={n} overlap={cfg}: {us:9.2f} us/step  {us * 65536 / n:7.2f} us per 65536 envs  {n / us * 1e6:.4g} env-steps/s",
              flush=True)
