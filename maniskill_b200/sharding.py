"""Env sharding across GPUs (one process per GPU).

Environments are independent: the controller step has NO collective.  Each rank owns a
contiguous block of env rows (its own PhysX-layout buffers); the only communication is
an ``all_gather`` of a tiny per-shard statistics vector per reporting interval (NCCL over
NVLink on GPUs, gloo in the CPU tests).  The reference has no multi-GPU code at all
(SURVEY K6); this is the env-sharded data parallelism north_star asks for.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Dict, List, Tuple

import torch
import torch.distributed as dist

STAT_FIELDS = ("solves", "elapsed_ms", "err_sum", "err_max", "iterations")


def shard_range(n_envs: int, rank: int, world_size: int) -> Tuple[int, int]:
    """[lo, hi) rows of rank ``rank``; sizes differ by at most one, order preserved."""
    assert 0 <= rank < world_size and n_envs >= 0
    base, rem = divmod(n_envs, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def env_world() -> Tuple[int, int, int]:
    """(rank, local_rank, world_size) from the torchrun environment (defaults: single process)."""
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def init_distributed(backend: str = "nccl") -> Tuple[int, int, int]:
    rank, local_rank, world = env_world()
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        kw = {}
        if backend == "nccl":
            # keep NCCL's banner off stdout (bench.py prints exactly one JSON line there)
            # (NCCL writes its debug output, version banner included, to stdout unless told otherwise)
            os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
            kw["device_id"] = torch.device("cuda", local_rank)
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
    return rank, local_rank, world


@dataclass
class ShardStats:
    solves: float = 0.0
    elapsed_ms: float = 0.0
    err_sum: float = 0.0
    err_max: float = 0.0
    iterations: float = 0.0

    def tensor(self, device) -> torch.Tensor:
        return torch.tensor([getattr(self, f) for f in STAT_FIELDS], dtype=torch.float64, device=device)


def gather_stats(stats: ShardStats, device="cpu") -> List[ShardStats]:
    """all_gather of the per-shard stats vector (40 B / rank)."""
    t = stats.tensor(device)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [stats]
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return [ShardStats(*[float(v) for v in o.tolist()]) for o in out]


def reduce_stats(all_stats: List[ShardStats]) -> Dict[str, float]:
    """whole-job aggregate: total solves, MAX elapsed over ranks, error summary."""
    solves = sum(s.solves for s in all_stats)
    elapsed = max(s.elapsed_ms for s in all_stats)
    return {
        "solves": solves,
        "elapsed_ms": elapsed,
        "solves_per_s": solves / (elapsed * 1e-3) if elapsed > 0 else 0.0,
        "err_mean": sum(s.err_sum for s in all_stats) / max(solves, 1.0),
        "err_max": max(s.err_max for s in all_stats),
        "iterations_mean": sum(s.iterations for s in all_stats) / max(solves, 1.0),
    }


def shutdown_distributed() -> None:
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()
