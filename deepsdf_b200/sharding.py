"""Host-side sharding rules of the multi-GPU path (one process per GPU, torch.distributed):

* training shards by SCENE: rank r owns a contiguous block of scenes (their lat_vecs rows, Adam moments and samples);
  every rank normalises its loss by the GLOBAL row count (train_deep_sdf.py:519), so the only exchange is a plain SUM
  all-reduce of the flat decoder gradient - no latent ever crosses a link;
* dropout streams are keyed by the global row id: rank r's rows start at ``row_offset``;
* grid evaluation shards by z-slab (contiguous range of the slowest index of create_mesh's ordering, mesh.py:45-47);
* the latent optimisation shards by shape - replicas only, no collective.

Backend-agnostic (NCCL on the GPUs, gloo in the CPU tests).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch


def block_partition(n: int, rank: int, world_size: int) -> Tuple[int, int]:
    """Contiguous range [i0, i1) of n items owned by ``rank``; sizes differ by at most one, earlier ranks get the extra."""
    if not 0 <= rank < world_size:
        raise ValueError(f"rank {rank} outside world of {world_size}")
    base, rem = divmod(n, world_size)
    i0 = rank * base + min(rank, rem)
    return i0, i0 + base + (1 if rank < rem else 0)


def partition_scenes(num_scenes: int, rank: int, world_size: int) -> range:
    return range(*block_partition(num_scenes, rank, world_size))


def slab_for_rank(N: int, rank: int, world_size: int) -> Tuple[int, int]:
    """z-slab [i0, i1) of the slowest grid index owned by ``rank``."""
    return block_partition(N, rank, world_size)


def row_offset(rank: int, rows_per_rank: int) -> int:
    """Global row id of this rank's local row 0 (weak scaling: every rank brings ``rows_per_rank`` rows per step)."""
    return rank * rows_per_rank


def allreduce_sum_(flat: torch.Tensor, group=None, world_size: Optional[int] = None) -> torch.Tensor:
    """In-place SUM all-reduce of the flat decoder gradient (replaces nn.DataParallel's gather, train_deep_sdf.py:353)."""
    import torch.distributed as dist
    ws = world_size if world_size is not None else (dist.get_world_size(group) if dist.is_initialized() else 1)
    if ws > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    return flat


def max_over_ranks(value: float, device, group=None) -> float:
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return value
    t = torch.tensor([value], device=device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())
