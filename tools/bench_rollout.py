#!/usr/bin/env python
"""BASELINE.json configs[4]: PPO rollout collection (train_test/config.py:3-15 shapes) against the GPU VecEnv with
device-resident observations / actions.  Prints rollout env-steps/s with the policy forward included and excluded."""
import argparse, json, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gym_usv_b200 as usv
from gym_usv_b200.rollout import DeviceRolloutBuffer, MlpActorCritic, collect_rollouts

ap = argparse.ArgumentParser()
ap.add_argument("--env", default="usv-simple")
ap.add_argument("--envs", type=int, default=4096)
ap.add_argument("--n-steps", type=int, default=2048)
ap.add_argument("--stack", type=int, default=5)
args = ap.parse_args()
dev = torch.device("cuda", 0)
base = usv.UsvGpuVecEnv(args.envs, device=0, seed=0, env_id=args.env, obstacles_f32=True)
env = usv.UsvGpuVecFrameStack(base, args.stack) if args.stack > 1 else base
d, a = env.observation_space.shape[0], base.act_dim
pol = MlpActorCritic(d, a).to(dev)
buf = DeviceRolloutBuffer.allocate(args.n_steps, args.envs, d, a, dev)
out = {"workload": f"PPO rollout, {args.env}, {args.envs} envs x n_steps {args.n_steps}, VecFrameStack({args.stack}), "
                   "MlpPolicy pi/vf [256, 256] ReLU (train_test/config.py:3-15), fp32 torch policy",
       "obs_dim": d}
for name, with_policy in (("with_policy", True), ("env_only", False)):
    obs = env.reset_tensor()
    starts = torch.ones(args.envs, device=dev)
    small = DeviceRolloutBuffer.allocate(8, args.envs, d, a, dev)
    collect_rollouts(env, pol, small, obs, starts, step_policy=with_policy)  # warm-up
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    obs, starts = collect_rollouts(env, pol, buf, obs, starts, step_policy=with_policy)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    out[name] = {"env_steps_per_sec": args.envs * args.n_steps / dt, "seconds": dt}
print(json.dumps(out))
