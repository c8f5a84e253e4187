"""Per-kernel device time of one env step (usvgpu_step_timed): python tools/time_step.py [env_id] [n_envs]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import gym_usv_b200 as usv  # noqa: E402

env_id = sys.argv[1] if len(sys.argv) > 1 else "usv-simple"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
envs = [usv.UsvGpuVecEnv(n, device=0, seed=s, env_id=env_id) for s in range(3)]
g = torch.Generator(device=envs[0].device).manual_seed(0)
a = (torch.rand((n, envs[0].act_dim), generator=g, device=envs[0].device) * 0.8 + 0.2).contiguous()
for e in envs:
    e.reset_tensor()
    for _ in range(20):
        e.step_tensor(a)
d = s = 0.0
K = 60
for k in range(K):
    dm, sm = envs[k % 3].step_tensor_timed(a)
    d += dm / K
    s += sm / K
print(f"{env_id} {n} envs: first kernel {d * 1e3:.1f} us, second {s * 1e3:.1f} us")
