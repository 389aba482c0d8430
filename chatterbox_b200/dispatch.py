"""Rank-level plumbing for N > 1: independent utterances shard across replicas, one process per
GPU, no collective on the data path.  `torch.distributed` is used for exactly two things: a barrier around the
timed region and a MAX reduction of device-measured times (NCCL on GPUs; gloo in the CPU tests)."""
from __future__ import annotations

from typing import List, Sequence

import torch
import torch.distributed as dist


def shard(n_items: int, rank: int, world: int) -> List[int]:
    """Utterance k runs on replica k mod world (`server.py` queues requests; each GPU is a full replica)."""
    assert 0 <= rank < world
    return list(range(rank, n_items, world))


def barrier(world: int) -> None:
    if world > 1:
        dist.barrier()


def max_over_ranks(values: Sequence[float], device, world: int) -> List[float]:
    """Element-wise maximum over ranks of device-measured durations (never a wall clock)."""
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.tolist()


def whole_job_rate(units_per_rank: float, world: int, ms: float) -> float:
    """Weak scaling: every rank processed `units_per_rank` units; the job took the slowest rank's `ms`."""
    return units_per_rank * world / (ms * 1e-3)
