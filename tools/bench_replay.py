"""Off-policy (SAC) data path on the device: env step + VecFrameStack(5) + ReplayBuffer.add, and ReplayBuffer.sample.

python tools/bench_replay.py [--envs 4096] [--steps 200] [--stack 5] [--buffer-size 400000]
One JSON line per run: env-steps/s of collect_transitions with uniform actions (the learning_starts phase of
train_test/config.py:17-37), the add kernel's HBM rate, and the latency of one 256-sample batch.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import gym_usv_b200 as usv
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=4096)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--stack", type=int, default=5)
    ap.add_argument("--buffer-size", type=int, default=400000)
    ap.add_argument("--env", default="usv-simple")
    args = ap.parse_args()
    venv = usv.UsvGpuVecEnv(args.envs, device=0, seed=0, env_id=args.env)
    env = usv.UsvGpuVecFrameStack(venv, n_stack=args.stack) if args.stack > 1 else venv
    d = venv.obs_dim * max(args.stack, 1)
    buf = usv.DeviceReplayBuffer(max(args.buffer_size, 4 * args.envs), args.envs, d, venv.act_dim, venv.device)
    g = torch.Generator(device=venv.device).manual_seed(0)
    obs = env.reset_tensor()
    obs = usv.collect_transitions(env, None, buf, obs, 20, generator=g)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    obs = usv.collect_transitions(env, None, buf, obs, args.steps, generator=g)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    # the add kernel alone
    hold = obs.clone()
    a = torch.rand((args.envs, venv.act_dim), device=venv.device)
    z8 = torch.zeros(args.envs, dtype=torch.uint8, device=venv.device)
    r = torch.zeros(args.envs, device=venv.device)
    for _ in range(3):
        buf.add(hold, obs, a, r, z8, z8, None)
    e0.record()
    for _ in range(50):
        buf.add(hold, obs, a, r, z8, z8, None)
    e1.record()
    torch.cuda.synchronize()
    add_ms = e0.elapsed_time(e1) / 50
    add_bytes = args.envs * (4 * d * 4 + 2 * venv.act_dim * 4 + 8 + 4)  # 2 rows read + 2 written, action, reward, flags
    for _ in range(3):
        buf.sample(256, generator=g)
    e0.record()
    for _ in range(100):
        buf.sample(256, generator=g)
    e1.record()
    torch.cuda.synchronize()
    sample_us = e0.elapsed_time(e1) * 10
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:  # noqa: BLE001
        peak = 6650.0
    print(json.dumps({"metric": "env_steps_per_sec", "what": "collect_transitions (uniform actions) + replay add",
                      "env": args.env, "envs": args.envs, "stack": args.stack, "obs_dim": d, "steps": args.steps,
                      "value": args.envs * args.steps / (ms * 1e-3), "ms_per_step": ms / args.steps,
                      "buffer_slots_per_env": buf.buffer_size, "buffer_gb": 2 * buf.observations.numel() * 4 / 1e9,
                      "replay_add": {"us": add_ms * 1e3, "bytes": add_bytes, "gbs": add_bytes / (add_ms * 1e-3) / 1e9,
                                     "frac_of_hbm_peak": add_bytes / (add_ms * 1e-3) / 1e9 / peak},
                      "sample_256_us": sample_us}))


if __name__ == "__main__":
    main()
