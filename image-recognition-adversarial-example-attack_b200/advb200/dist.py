"""Multi-GPU plumbing: the image batch is embarrassingly parallel, so each rank attacks a contiguous
shard and the only exchange is the int64[7] counter vector (defense_experiments.py:809-832 sums) —
one all-reduce per (attack, eps) — plus a broadcast of the calibrated detector threshold.
NCCL over NVLink on GPUs; the same code runs over gloo for the CPU tests."""
from __future__ import annotations

import os
from typing import Tuple

import torch
import torch.distributed as dist


def env_rank_world() -> Tuple[int, int, int]:
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def init(backend: str | None = None) -> Tuple[int, int, int]:
    """Initialise torch.distributed from the torchrun environment (no-op for a single process)."""
    rank, world, local = env_rank_world()
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend)
    return rank, world, local


def shard_range(n_total: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous shard [lo, hi) of rank; the first n_total % world ranks take one extra image."""
    base, rem = divmod(n_total, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def allreduce_counters(counters: torch.Tensor) -> torch.Tensor:
    """Sum the per-rank int64 counter vector over all ranks (in place); identity for one process."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(counters, op=dist.ReduceOp.SUM)
    return counters


def broadcast_scalar(value: float, src: int = 0, device=None) -> float:
    """Rank `src`'s float to every rank (the detector threshold is calibrated on rank 0's images)."""
    if not (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1):
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.broadcast(t, src=src)
    return float(t.item())


def max_over_ranks(value: float, device=None) -> float:
    if not (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1):
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier() -> None:
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()


def rates_from_counters(c) -> dict:
    """The five rates the reference prints per (attack, eps) (defense_experiments.py:841-850)."""
    c = [int(v) for v in c]
    count = max(1, c[6])
    return {
        "attack_success": c[1] / count,
        "preproc_defense_acc": c[2] / count,
        "detector_clean_pass_rate": 1.0 - c[3] / count,
        "detector_adv_flag_rate": c[4] / count,
        "detector_attack_success": c[5] / count,
    }


def shutdown() -> None:
    """Tear the process group down (quietly a no-op for single-process runs)."""
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()
