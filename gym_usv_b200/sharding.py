"""Multi-GPU layout of the env batch: contiguous env-index ranges, one process per GPU.

Envs are independent (each reference env object owns its arrays: gym_usv/envs/simple_env.py:38-54,
gym_usv/control/usv_asmc.py:43-49), so the step path has NO exchange step and no collective.  The
only cross-rank traffic is the optional reduction of the eight episode-statistics scalars that SB3's
Monitor / wandb logging would aggregate (train_test/sb3_train_vec.py:60-65), done with one
``torch.distributed.all_reduce`` (NCCL on GPUs, gloo in the CPU tests).

The reset RNG is keyed by the GLOBAL env id (csrc/usv_philox.cuh), so the trajectories of the job do
not depend on the number of ranks: rank r simply passes ``global_env_offset = shard_range(...)[0]``.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch

STAT_KEYS = ("return_sum", "length_sum", "episodes", "terminated", "out_of_bounds", "time_limit", "env_steps",
             "nonfinite")


def shard_range(n_total: int, rank: int, world_size: int) -> Tuple[int, int]:
    """[lo, hi) of the envs owned by ``rank``: contiguous, sizes differ by at most one, earlier ranks
    take the remainder."""
    if world_size <= 0 or not 0 <= rank < world_size:
        raise ValueError(f"bad rank/world_size {rank}/{world_size}")
    if n_total < 0:
        raise ValueError("n_total must be >= 0")
    base, rem = divmod(int(n_total), int(world_size))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def owner_of(env_id: int, n_total: int, world_size: int) -> int:
    """Rank that owns global env ``env_id`` under shard_range()."""
    base, rem = divmod(int(n_total), int(world_size))
    cut = rem * (base + 1)
    if env_id < cut:
        return env_id // (base + 1)
    return rem + (env_id - cut) // max(base, 1)


def reduce_episode_stats(stats: Dict[str, float], device: Optional[torch.device] = None, group=None) -> Dict[str, float]:
    """Sum the per-rank UsvStats dict over all ranks (float64 carrier: counters are < 2**53).
    With no initialised process group this is the identity."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return dict(stats)
    t = torch.tensor([float(stats[k]) for k in STAT_KEYS], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    vals = t.cpu().tolist()
    out = {}
    for k, v in zip(STAT_KEYS, vals):
        out[k] = v if k in ("return_sum", "length_sum") else int(round(v))
    return out


def make_sharded_env(n_total: int, rank: int, world_size: int, local_device: int, seed: int = 0, **kwargs):
    """The rank's slice of an ``n_total``-env job as a UsvGpuVecEnv (BASELINE.json config 4)."""
    from .vec_env import UsvGpuVecEnv
    lo, hi = shard_range(n_total, rank, world_size)
    return UsvGpuVecEnv(hi - lo, device=local_device, seed=seed, global_env_offset=lo, **kwargs)
