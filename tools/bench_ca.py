"""'usv-asmc-ca-v0' (collision-avoidance env, one ca_kernel launch per VecEnv.step) timed on the device.

python tools/bench_ca.py [--envs 262144] [--steps 200] [--warmup 50] [--cpu-envs 8192] [--cpu-steps 20]
One JSON line: env-steps/s with the state resident in HBM (CUDA events around `steps` launches, rotating action
slabs, episodes pre-rolled so that in-kernel resets are part of the steady state), the kernel's algorithmic bytes
against the HBM peak, and the C restatement of the same env on the host cores as the CPU baseline.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

# per env-step: read 22 dyn rows + 31 ca rows (f64), n_obs, steps (i32), return (f64), action (2 f32);
# write 22 dyn rows, previous action (2 f64), steps, return, observation (23 f32), reward, 2 flag bytes
CA_BYTES = (22 * 8 + 31 * 8 + 4 + 4 + 8 + 8) + (22 * 8 + 2 * 8 + 4 + 8 + 23 * 4 + 4 + 2)


def main():
    import numpy as np
    import torch
    import gym_usv_b200 as usv
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=262144)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--cpu-envs", type=int, default=8192)
    ap.add_argument("--cpu-steps", type=int, default=20)
    args = ap.parse_args()
    env = usv.UsvGpuVecEnv(args.envs, device=0, seed=0, env_id="usv-asmc-ca-v0")
    dev = env.device
    gen = torch.Generator(device=dev).manual_seed(1)
    acts = (torch.rand((8, args.envs, 2), generator=gen, device=dev) * 2 - 1).contiguous()
    env.reset_tensor()
    for k in range(args.warmup):
        env.step_tensor(acts[k % 8])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(args.steps):
        env.step_tensor(acts[k % 8])
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    stats = env.episode_stats()
    value = args.envs / (ms * 1e-3)
    peak = None
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peak = float(json.load(f)["hbm_gbs"])
    except Exception:  # noqa: BLE001
        pass
    line = {"metric": "env_steps_per_sec", "env": "usv-asmc-ca-v0", "value": value, "unit": "env-steps/s", "n_gpus": 1,
            "envs": args.envs, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "dtype": "f64",
            "sub_steps_per_env_step": 10, "gpu_launches": args.steps,
            "episodes_in_run": stats["episodes"], "nonfinite": stats["nonfinite"],
            "roofline": {"bound": "fp64 pipe (10 sub-steps + 16 x n_obs ray tests per env-step); HBM figure for scale",
                         "algorithmic_bytes_per_env_step": CA_BYTES, "achieved": CA_BYTES * value / 1e9, "unit": "GB/s",
                         "peak": peak, "frac": (CA_BYTES * value / 1e9 / peak) if peak else None}}
    if args.cpu_envs > 0:
        import usv_oracle as uo
        threads = len(os.sched_getaffinity(0))
        ov = uo.OracleCaVecEnv(args.cpu_envs, seed=0, n_threads=threads)
        a = np.random.default_rng(0).uniform(-1, 1, (8, args.cpu_envs, 2)).astype(np.float32)
        ov.reset()
        for k in range(3):
            ov.step(a[k % 8])
        t0 = time.perf_counter()
        for k in range(args.cpu_steps):
            ov.step(a[k % 8])
        el = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": args.cpu_envs * args.cpu_steps / el, "unit": "env-steps/s", "cores": threads,
                                "kind": "port", "sample": f"{args.cpu_envs} lock-step envs x {args.cpu_steps} steps, OpenMP"}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
