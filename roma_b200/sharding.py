"""Pair sharding across GPUs: pairs are independent, so the path shards by pair index with no
data-path collective.  The only communication is the bookkeeping below (backend-agnostic: NCCL on the GPU box,
gloo in the CPU tests)."""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.distributed as dist


def shard_range(n_pairs: int, world: int, rank: int) -> range:
    """Contiguous shard of pair indices for `rank`: pair i -> rank i // ceil(n/world) (BASELINE.json configs[2])."""
    if not (0 <= rank < world) or n_pairs < 0:
        raise ValueError(f"bad shard request n_pairs={n_pairs} world={world} rank={rank}")
    per = (n_pairs + world - 1) // world
    lo = min(rank * per, n_pairs)
    return range(lo, min(lo + per, n_pairs))


def aggregate(local_pairs: int, local_ms: float, device: Optional[torch.device] = None) -> Tuple[int, float]:
    """(total pairs processed by all ranks, max over ranks of the timed milliseconds)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return local_pairs, local_ms
    dev = device or torch.device("cpu")
    t = torch.tensor([float(local_ms)], dtype=torch.float64, device=dev)
    n = torch.tensor([int(local_pairs)], dtype=torch.int64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(n, op=dist.ReduceOp.SUM)
    return int(n.item()), float(t.item())
