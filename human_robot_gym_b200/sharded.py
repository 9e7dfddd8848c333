"""Env-wise sharding of a batch of shields over the GPUs of one box (one process per GPU, torch.distributed).

Every env's shield state is private (the reference keeps one SafetyShield per robot per env, human_env.py:1395-1413), so
the data path needs NO collective: rank r owns the contiguous env block shard_range(N, r, world). The only exchange is
the reduction of the per-step verify statistics (HrgStats): one all_reduce(SUM) of nine int64 counters and one
all_reduce(MIN) of the smallest capsule distance, issued off the tick critical path (NCCL over NVLink on the GPUs;
gloo in the CPU tests).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.distributed as dist

STAT_FIELDS = ("n_steps", "n_unsafe", "n_branch0", "n_branch1", "n_branch2", "n_replan", "n_replan_failed",
               "n_near_threshold", "ltt_len_sum")  # then min_dist (f64 bit pattern) — struct HrgStats, include/hrg_shield.h
TAIL_FIELDS = ("n_ltt_truncated", "n_pfl_rejected")  # after min_dist


def shard_range(n_total: int, rank: int, world: int) -> Tuple[int, int]:
    """[lo, hi) of rank's contiguous env block; the first n_total % world ranks get one extra env."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    base, extra = divmod(int(n_total), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard(x, rank: int, world: int, n_total: Optional[int] = None):
    """The rows of a global [N, ...] array / tensor that belong to `rank`."""
    n = int(n_total if n_total is not None else x.shape[0])
    lo, hi = shard_range(n, rank, world)
    return x[lo:hi]


def reduce_stats(stats: torch.Tensor, group=None) -> dict:
    """All-reduce one raw HrgStats tensor (int64 words, word 9 = min_dist bits) over the ranks -> dict.

    Works on whatever device the tensor lives on (cuda with the nccl backend, cpu with gloo). Without an initialised
    process group the local values are returned unchanged."""
    counts = torch.cat([stats[:9], stats[10:10 + len(TAIL_FIELDS)]])
    mind = stats[9:10].clone().view(torch.float64)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(mind, op=dist.ReduceOp.MIN, group=group)
    c = counts.cpu().tolist()
    out = dict(zip(STAT_FIELDS + TAIL_FIELDS[:max(0, len(c) - 9)], (int(v) for v in c)))
    out["min_dist"] = float(mind.item())
    out["n_branch"] = [out.pop("n_branch0"), out.pop("n_branch1"), out.pop("n_branch2")]
    return out


def max_over_ranks(ms: float, device, group=None) -> float:
    """Device-timed duration -> max over ranks (how every multi-GPU number is reported)."""
    t = torch.tensor([ms], dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


class ShardedShield:
    """This rank's block of a global batch of `n_total` shields: a BatchedSafetyShield over shard_range(...)."""

    def __init__(self, n_total: int, sample_time: float, trajectory_config_file: str, robot_config_file: str,
                 mocap_config_file: str, base_pose=None, init_qpos=None, rank: Optional[int] = None,
                 world: Optional[int] = None, device=None, **kw):
        from .batched import BatchedSafetyShield
        self.rank = dist.get_rank() if rank is None else rank
        self.world = dist.get_world_size() if world is None else world
        self.n_total = int(n_total)
        self.lo, self.hi = shard_range(n_total, self.rank, self.world)
        bp = None if base_pose is None else torch.as_tensor(base_pose)
        iq = None if init_qpos is None else torch.as_tensor(init_qpos)
        if bp is not None and bp.ndim == 2:
            bp = bp[self.lo:self.hi]
        if iq is not None and iq.ndim == 2:
            iq = iq[self.lo:self.hi]
        if "shield_type" in kw:  # a per-env ShieldType array is sharded like base_pose / init_qpos
            import numpy as np
            ty = np.asarray(kw["shield_type"])
            if ty.ndim == 1:
                kw = dict(kw, shield_type=ty[self.lo:self.hi])
        self.local = BatchedSafetyShield(self.hi - self.lo, sample_time, trajectory_config_file, robot_config_file,
                                         mocap_config_file, base_pose=None if bp is None else bp.numpy(),
                                         init_qpos=None if iq is None else iq.numpy(), device=device, **kw)

    def rows(self, x):
        return x[self.lo:self.hi]

    def global_stats(self, clear=False, group=None) -> dict:
        return reduce_stats(self.local.stats_tensor(clear=clear), group)
