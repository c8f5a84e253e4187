"""torchrun worker of tests/test_multi_gpu.py: one rank per GPU, envs sharded by contiguous index range.

Every rank steps ITS slice of an N-env job on ITS device for T steps; rank 0 also runs the whole job on cuda:0.  The
slices, gathered over NCCL, must equal the matching rows of the single-GPU run bit for bit (the reset RNG is keyed by
the global env id, csrc/usv_philox.cuh), and reduce_episode_stats over NCCL must equal the sum of the per-rank structs.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gym_usv_b200 as usv  # noqa: E402
from gym_usv_b200.sharding import STAT_KEYS, reduce_episode_stats, shard_range  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    n_total, steps, seed = int(sys.argv[1]), int(sys.argv[2]), 21
    env_id = sys.argv[3] if len(sys.argv) > 3 else "usv-asmc-simple"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    lo, hi = shard_range(n_total, rank, world)
    kw = dict(seed=seed, env_id=env_id, max_episode_steps=25)  # a short TimeLimit puts the in-kernel reset into the run
    env = usv.UsvGpuVecEnv(hi - lo, device=local, global_env_offset=lo, **kw)
    act_dim = env.act_dim
    acts = np.random.default_rng(5).uniform(-1 if act_dim == 1 else [0.2, -1.0], 1 if act_dim == 1 else [1.0, 1.0],
                                            size=(steps, n_total, act_dim)).astype(np.float32)
    full = usv.UsvGpuVecEnv(n_total, device=0, **kw) if rank == 0 else None
    o = env.reset_tensor().clone()
    if rank == 0:
        o_full = full.reset_tensor().clone()
    bad = 0
    for t in range(steps + 1):
        if t > 0:
            o, r, te, tr = [x.clone() for x in env.step_tensor(torch.tensor(acts[t - 1, lo:hi], device=dev))]
        else:
            r = torch.zeros(hi - lo, device=dev)
            te = tr = torch.zeros(hi - lo, dtype=torch.uint8, device=dev)
        # variable-size gather over NCCL: pad every shard to the largest one
        m = (n_total + world - 1) // world
        pack = torch.zeros((m, env.obs_dim + 3), dtype=torch.float32, device=dev)
        pack[:hi - lo, :env.obs_dim] = o
        pack[:hi - lo, env.obs_dim] = r
        pack[:hi - lo, env.obs_dim + 1] = te.float()
        pack[:hi - lo, env.obs_dim + 2] = tr.float()
        parts = [torch.empty_like(pack) for _ in range(world)] if rank == 0 else None
        dist.gather(pack, parts, dst=0)
        if rank == 0:
            if t > 0:
                o_full, r_full, te_full, tr_full = [x.clone() for x in full.step_tensor(torch.tensor(acts[t - 1], device=dev))]
            else:
                r_full = torch.zeros(n_total, device=dev)
                te_full = tr_full = torch.zeros(n_total, dtype=torch.uint8, device=dev)
            for q in range(world):
                a, b = shard_range(n_total, q, world)
                got = parts[q][:b - a]
                same = (torch.equal(got[:, :env.obs_dim], o_full[a:b]) and torch.equal(got[:, env.obs_dim], r_full[a:b])
                        and torch.equal(got[:, env.obs_dim + 1], te_full[a:b].float())
                        and torch.equal(got[:, env.obs_dim + 2], tr_full[a:b].float()))
                if not same:
                    bad += 1
                    print(f"MISMATCH step {t} rank {q}", flush=True)
    mine = env.episode_stats()
    total = reduce_episode_stats(mine, device=dev)        # NCCL all-reduce of the eight scalars
    every = [None] * world
    dist.all_gather_object(every, mine)
    want = {k: sum(e[k] for e in every) for k in STAT_KEYS}
    stats_ok = all(abs(total[k] - want[k]) <= 1e-9 * max(1.0, abs(want[k])) for k in STAT_KEYS)
    if rank == 0:
        whole = full.episode_stats()
        whole_ok = all(abs(whole[k] - want[k]) <= 1e-6 * max(1.0, abs(want[k])) for k in STAT_KEYS)
        print(f"ranks {world} envs {n_total} steps {steps} env {env_id}: mismatching (step, rank) pairs {bad}, episodes "
              f"{total['episodes']} (single GPU {whole['episodes']}), NCCL stats reduce {'ok' if stats_ok else 'BAD'}, "
              f"sum of shards == single-GPU stats {'ok' if whole_ok else 'BAD'}", flush=True)
        if bad == 0 and stats_ok and whole_ok and total["episodes"] > 0:
            print("MULTI_GPU_OK", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
