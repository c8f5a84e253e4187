"""Launches the kernels beside the headline pair once each at their bench sizes, for an ncu capture:
legacy_step_kernel / legacy64_step_kernel<1|2> (1 M envs), render_kernel (256 frames of 512 x 512),
frame_push_kernel + replay_add_kernel (65 536 envs, VecFrameStack(5)), kinematic_kernel."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import math
    import torch
    import gym_usv_b200 as usv
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(0)
    for env_id in ("usv-asmc-v0", "usv-asmc-ye-int-v0", "usv-pid-v0"):
        env = usv.UsvGpuVecEnv(1 << 20, device=0, seed=0, env_id=env_id)
        env.reset_tensor()
        for _ in range(6):
            a = (torch.rand((1 << 20, 1), generator=g, device=dev) - 0.5) * math.pi
            env.step_tensor(a.contiguous())
        torch.cuda.synchronize()
        env.close()
    venv = usv.UsvGpuVecEnv(65536, device=0, seed=0, env_id="usv-simple")
    env = usv.UsvGpuVecFrameStack(venv, n_stack=5)
    buf = usv.DeviceReplayBuffer(8 * 65536, 65536, 5 * venv.obs_dim, venv.act_dim, dev)
    obs = env.reset_tensor()
    obs = usv.collect_transitions(env, None, buf, obs, 6, generator=g)
    out = venv.render_tensor(list(range(256)))
    out = venv.render_tensor(list(range(256)))
    torch.cuda.synchronize()
    print("ok", tuple(out.shape), buf.pos)


if __name__ == "__main__":
    main()
