"""Multi-GPU partitioning of independent stereo sequences.

Sequences (camera streams) are independent: all state is per sequence
(src/feature_tracker.hpp:57-62, src/bundle_adjuster.hpp:134-149), so the path shards with NO
data-path collective. One process per GPU owns a contiguous block of global sequence ids; the
only cross-rank step is the max-reduction of the timed region (and a frame-count sum), done
with torch.distributed (nccl on GPUs, gloo in the CPU tests).
"""
from __future__ import annotations

from typing import List


def shard_sequences(n_total: int, world: int, rank: int) -> List[int]:
    """Global sequence ids owned by `rank` when n_total sequences are split over `world` ranks
    (strong scaling: block partition, remainder spread over the first ranks)."""
    if world <= 0 or not (0 <= rank < world) or n_total < 0:
        raise ValueError("bad shard request")
    base, rem = divmod(n_total, world)
    start = rank * base + min(rank, rem)
    return list(range(start, start + base + (1 if rank < rem else 0)))


def weak_shard(seqs_per_gpu: int, world: int, rank: int) -> List[int]:
    """Weak scaling: every rank owns `seqs_per_gpu` sequences; ids are globally unique."""
    if seqs_per_gpu < 0 or world <= 0 or not (0 <= rank < world):
        raise ValueError("bad shard request")
    return list(range(rank * seqs_per_gpu, (rank + 1) * seqs_per_gpu))


def owner_of(seq: int, n_total: int, world: int) -> int:
    """Inverse of shard_sequences."""
    base, rem = divmod(n_total, world)
    edge = rem * (base + 1)
    if seq < edge:
        return seq // (base + 1)
    return rem + (seq - edge) // base if base else world - 1


def reduce_time_and_frames(ms: float, frames: int, dist=None, device=None):
    """(max over ranks of the timed region, total frames over ranks)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return ms, frames
    import torch
    t = torch.tensor([ms], dtype=torch.float64, device=device)
    f = torch.tensor([frames], dtype=torch.int64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(f, op=dist.ReduceOp.SUM)
    return float(t.item()), int(f.item())
