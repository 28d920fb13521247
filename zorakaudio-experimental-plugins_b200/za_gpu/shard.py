"""Multi-GPU plumbing: streams are independent, so a job shards as contiguous stream ranges over the
ranks (stream s -> rank s*G/n_streams) with NO collective on the data path.
torch.distributed is used only for the barrier, the max-over-ranks timing and the optional gather of
per-stream checksums to rank 0."""
from __future__ import annotations

import numpy as np


def shard_range(n_streams: int, rank: int, world: int) -> tuple[int, int]:
    """[first, last) streams of `rank`: contiguous, disjoint, covering, sizes differ by at most one."""
    return (n_streams * rank) // world, (n_streams * (rank + 1)) // world


def stream_checksums(y: np.ndarray) -> np.ndarray:
    """Per-stream checksum + energy of a rendered block [streams][ch][frames] (float32): what a sharded bounce
    reports instead of shipping audio back."""
    b = np.ascontiguousarray(y).view(np.uint32).reshape(y.shape[0], -1).astype(np.uint64)
    w = (np.arange(b.shape[1], dtype=np.uint64) * np.uint64(2654435761) + np.uint64(1)) & np.uint64(0xFFFFFFFF)
    chk = (b * w).sum(axis=1, dtype=np.uint64)
    en = (y.astype(np.float64) ** 2).sum(axis=(1, 2))
    return np.stack([chk.astype(np.float64) % 2 ** 52, en], axis=1)


def gather_to_rank0(dist, local: np.ndarray, n_streams: int, world: int):
    """All ranks' per-stream rows -> rank 0, in stream order (one small all_gather, outside any timed region)."""
    import torch
    counts = [shard_range(n_streams, r, world)[1] - shard_range(n_streams, r, world)[0] for r in range(world)]
    mx = max(counts)
    pad = np.zeros((mx, local.shape[1]), np.float64)
    pad[:local.shape[0]] = local
    bufs = [torch.zeros((mx, local.shape[1]), dtype=torch.float64) for _ in range(world)]
    dist.all_gather(bufs, torch.from_numpy(pad))
    return np.concatenate([b.numpy()[:c] for b, c in zip(bufs, counts)], axis=0)


def max_over_ranks(dist, seconds: float) -> float:
    import torch
    t = torch.tensor([seconds], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
