"""Corpus-scale offline extraction: sessions sharded over ranks (one process per GPU), baseline statistics combined with
ONE small collective (reference loops: prepare_corpus.py:179-236, baseline_offline.py:37-60)."""
from __future__ import annotations

from typing import Dict, Iterable, List, Optional, Sequence


def shard_sessions(n_sessions: int, rank: int, world: int) -> List[int]:
    """Round-robin assignment of session indices to ranks (no data-path exchange: sessions are independent)."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world {world}")
    return list(range(rank, n_sessions, world))


def gather_accumulators(acc, group=None):
    """All-gather a (k, 3, C) or (3, C) block of (n, mean, M2) accumulators -> (world * k, 3, C), rank-major.

    Works on CUDA tensors over NCCL (NVLink) and on CPU tensors over gloo (host-logic tests).  The merge itself is
    `ops.ChannelStats.merge` (a CUDA kernel, fixed order -> bit-identical mu/sigma on every rank)."""
    import torch
    import torch.distributed as dist

    a = acc.reshape(-1, 3, acc.shape[-1]).contiguous()
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return a
    world = dist.get_world_size(group)
    out = torch.empty((world * a.shape[0],) + tuple(a.shape[1:]), dtype=a.dtype, device=a.device)
    dist.all_gather_into_tensor(out, a, group=group)
    return out


def day_slots(days: Sequence[int]) -> Dict[int, int]:
    """Stable slot index per distinct day label (every rank must derive the same table)."""
    return {d: i for i, d in enumerate(sorted(set(days)))}
