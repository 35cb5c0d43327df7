"""Multi-GPU sharding of a batch of independent environments.

Environments never interact, so the data path has NO collective: rank r owns a contiguous range
of global env ids and keys its Philox streams with them (results do not depend on the number
of GPUs).  The only exchange is the all-reduce of the fp64[8] episode-statistics vector.
"""
from __future__ import annotations

from typing import Tuple

import torch


def shard_range(total_envs: int, world_size: int, rank: int) -> Tuple[int, int]:
    """(first global env id, number of local envs) of `rank`: contiguous, sizes differ by <= 1."""
    if not (0 <= rank < world_size):
        raise ValueError("rank out of range")
    base, rem = divmod(int(total_envs), int(world_size))
    n_local = base + (1 if rank < rem else 0)
    first = rank * base + min(rank, rem)
    return first, n_local


def all_reduce_stats(stats: torch.Tensor) -> torch.Tensor:
    """SUM all-reduce of the statistics vector over the default process group (NCCL for CUDA
    tensors, gloo for CPU tensors); identity when torch.distributed is not initialised."""
    if torch.distributed.is_available() and torch.distributed.is_initialized():
        torch.distributed.all_reduce(stats, op=torch.distributed.ReduceOp.SUM)
    return stats


def max_over_ranks(seconds: float, device=None) -> float:
    """MAX over ranks of a device-timed duration (multi-GPU numbers are the slowest rank's)."""
    t = torch.tensor([seconds], dtype=torch.float64, device=device)
    if torch.distributed.is_available() and torch.distributed.is_initialized():
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    return float(t.item())
