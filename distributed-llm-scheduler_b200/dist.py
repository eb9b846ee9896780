"""Multi-GPU: instance sharding + ONE all-gather of the per-instance result records.

The path is embarrassingly parallel over DAG instances: rank g of G takes the contiguous range
[g*B/G, (g+1)*B/G), runs it with no data-path collective, and the ranks exchange only the fixed-size
dls_summary_t records (48 bytes per instance and algorithm: makespan, peak memory, counters). Per-task outputs stay sharded.
"""
from __future__ import annotations

from typing import Tuple

import numpy as np

from . import _native as N
from .lowering import DagBatch


def shard_bounds(B: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous instance range of `rank`: [rank*B/world, (rank+1)*B/world)."""
    return (rank * B) // world, ((rank + 1) * B) // world


def gather_summaries(summary, world: int):
    """all_gather of this rank's [A, B_local, 48]-byte record tensor -> [world, A, B_max, 48].

    Shards may differ by one instance; they are padded to the largest shard before the collective.
    Works on CUDA tensors under NCCL and on CPU tensors under gloo.
    """
    import torch
    import torch.distributed as dist
    if world == 1:
        return summary.unsqueeze(0)
    A, b_local = summary.shape[0], summary.shape[1]
    sizes = torch.tensor([b_local], dtype=torch.int64, device=summary.device)
    all_sizes = [torch.zeros_like(sizes) for _ in range(world)]
    # (8 bytes per rank, every call: a cached "shards are equal" would let ranks disagree on which collective comes next
    # as soon as one call has even shards and the next has not)
    dist.all_gather(all_sizes, sizes)
    b_max = int(torch.cat(all_sizes).max().item())  # (one device-to-host read, not one per rank)
    if b_max != b_local:
        pad = torch.zeros((A, b_max - b_local, summary.shape[2]), dtype=summary.dtype, device=summary.device)
        summary = torch.cat([summary, pad], 1)
    out = torch.empty((world,) + tuple(summary.shape), dtype=summary.dtype, device=summary.device)
    if summary.is_cuda:
        dist.all_gather_into_tensor(out, summary.contiguous())
    else:
        parts = [out[r] for r in range(world)]
        dist.all_gather(parts, summary.contiguous())
    return out


def merge_gathered(gathered, B: int, world: int) -> np.ndarray:
    """[world, A, B_max, 48] bytes -> structured numpy [A, B] in global instance order."""
    g = gathered.cpu().numpy()
    A = g.shape[1]
    out = np.zeros((A, B), np.dtype(N.SUMMARY_DTYPE))
    for r in range(world):
        lo, hi = shard_bounds(B, world, r)
        rec = np.ascontiguousarray(g[r, :, :hi - lo]).view(np.dtype(N.SUMMARY_DTYPE)).reshape(A, hi - lo)
        out[:, lo:hi] = rec
    return out


def schedule_sharded(batch: DagBatch, algo_mask: int = N.MASK_ALL, device=None):
    """Run this rank's shard of `batch` on its GPU and all-gather the per-instance records.

    Returns (local BatchResult, summaries of ALL instances as a structured numpy array [A, B])."""
    import torch.distributed as dist
    from .batch import schedule_batch
    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    lo, hi = shard_bounds(batch.B, world, rank)
    local = schedule_batch(batch.shard(lo, hi), algo_mask, device=device)
    gathered = gather_summaries(local.summary, world)
    return local, merge_gathered(gathered, batch.B, world)
