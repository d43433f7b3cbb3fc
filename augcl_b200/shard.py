"""Multi-GPU layout of the hot path: independent shards, no data-path collective.

The reference is single-GPU; its unit of parallelism is "one agent / seed per GPU" (scripts/*.sh:1,
script.py:28-48).  Here every rank owns one replay ring, one background-bank replica and one Philox
stream; nothing crosses NVLink on the hot path.  torch.distributed is used only around a measurement:
a barrier before/after the timed region and ONE scalar all_reduce(MAX) of the elapsed time after it.
These helpers hold that host logic so it can be tested on CPU with the gloo backend.
"""
from __future__ import annotations

import os
from dataclasses import dataclass

import torch
import torch.distributed as dist


@dataclass(frozen=True)
class ShardPlan:
    rank: int
    world: int
    local_rank: int
    capacity: int          # transitions in THIS rank's ring
    batch: int             # samples THIS rank gathers per step
    ring_seed: int         # synthetic-frame seed of this rank's ring
    draw_seed: int         # Philox key of this rank's index / offset / weight draws
    global_batch: int


def env_rank():
    """(rank, world, local_rank) from the torchrun environment (1-process defaults)."""
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def plan(capacity_per_gpu: int, batch_per_gpu: int, base_seed: int = 0, rank=None, world=None, local_rank=None) -> ShardPlan:
    """Weak scaling: per-GPU ring and batch are fixed, seeds are distinct per rank."""
    r, w, lr = env_rank()
    rank = r if rank is None else rank
    world = w if world is None else world
    local_rank = lr if local_rank is None else local_rank
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world of {world}")
    return ShardPlan(rank, world, local_rank, int(capacity_per_gpu), int(batch_per_gpu), ring_seed=1000 + base_seed + rank,
                     draw_seed=base_seed + rank, global_batch=int(batch_per_gpu) * world)


def core_slice(local_rank: int, local_world: int, cores) -> list:
    """The cores rank `local_rank` of `local_world` co-located ranks should run on: an equal, contiguous slice of the
    allowed set (at least one core each; with more ranks than cores they wrap).  Each rank drives its own launch loop;
    without pinning the scheduler migrates them onto each other and the slowest rank's launch cadence shows up as lost
    scaling although no byte crosses between GPUs."""
    cores = sorted(cores)
    n = len(cores)
    if n == 0 or local_world <= 0:
        return []
    per = max(1, n // local_world)
    lo = (local_rank * per) % n
    return cores[lo:lo + per] or [cores[lo]]


def pin_to_cores(local_rank=None, local_world=None):
    """Pin this process to its core slice (Linux).  Returns the cores, or None when affinity is not available."""
    if not hasattr(os, "sched_setaffinity"):
        return None
    lr = int(os.environ.get("LOCAL_RANK", "0")) if local_rank is None else local_rank
    lw = int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1"))) if local_world is None else local_world
    try:
        cores = core_slice(lr, lw, os.sched_getaffinity(0))
        if cores:
            os.sched_setaffinity(0, cores)
        return cores
    except OSError:
        return None


def gather_over_ranks(value: float, device=None) -> list:
    """Every rank's host scalar, in rank order (identity without a process group).  Off the data path."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [float(value)]
    t = torch.tensor([float(value)], dtype=torch.float64, device=device if device is not None else "cpu")
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return [float(x.item()) for x in out]


def barrier():
    if dist.is_available() and dist.is_initialized():
        dist.barrier()


def max_over_ranks(value: float, device=None) -> float:
    """MAX of a host scalar over all ranks (identity without a process group).  Off the data path."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def aggregate_rate(samples_per_rank: int, world: int, seconds_max: float) -> float:
    """Whole-job samples/s: every rank processed samples_per_rank in at most seconds_max."""
    return world * samples_per_rank / seconds_max
