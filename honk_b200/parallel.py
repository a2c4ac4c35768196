"""Multi-GPU plumbing: one process per GPU, contiguous clip shards, one gather of logits.

Clips are independent (BatchNorm uses running statistics in eval mode), so the hot path has
no data-path collective; the only exchange is the final all_gather of ``[B/R, n_labels]``
logits (48 B/clip) over NCCL/NVLink - or gloo in the CPU tests.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_range(n_clips, rank, world_size):
    """Contiguous ``[lo, hi)`` of rank's clips; the first ``n % R`` ranks take one extra."""
    base, rem = divmod(int(n_clips), int(world_size))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def gather_logits(local, n_total, group=None):
    """all_gather of ragged ``[n_r, L]`` shards into ``[n_total, L]`` on every rank, in shard order."""
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return local
    world = dist.get_world_size(group)
    L = local.shape[1]
    cap = (n_total + world - 1) // world
    buf = local.new_zeros((cap, L))
    buf[: local.shape[0]] = local
    out = local.new_empty((world * cap, L))
    dist.all_gather_into_tensor(out, buf, group=group)
    pieces = []
    for r in range(world):
        lo, hi = shard_range(n_total, r, world)
        pieces.append(out[r * cap: r * cap + (hi - lo)])
    return torch.cat(pieces, 0)
