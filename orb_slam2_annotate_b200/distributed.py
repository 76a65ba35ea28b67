"""Multi-GPU plumbing (one process per GPU, torch.distributed; NCCL on the GPU box, gloo in the
CPU tests).  The front end shards naturally, so there is no data-path collective:

  * extraction: frame i of a stream goes to rank i mod G (independent units, results stay with
    the frame's rank);
  * brute-force kNN-2 (config 5): QUERY rows are sharded contiguously, the train set is
    replicated, and the only exchange is one all_gather of the per-query results
    (idx, d1, d2) — 12 bytes per query.
"""
from __future__ import annotations

import numpy as np


def shard_range(n: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous shard [begin, end) of n rows for `rank`; sizes differ by at most one."""
    base, rem = divmod(n, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def frames_for_rank(nframes: int, rank: int, world: int) -> np.ndarray:
    """Frame i -> rank i mod G."""
    return np.arange(rank, nframes, world)


def knn2_sharded(knn2_fn, query: np.ndarray, train: np.ndarray, group=None, device=None):
    """Every rank computes knn2_fn(query[shard], train) -> (idx, d1, d2) for its query shard and
    the shards are all-gathered, so every rank returns the full (idx, d1, d2) of all queries.
    `knn2_fn` is the rank's matcher (ORBmatcher.knn2 on the GPU box)."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    m = len(query)
    b, e = shard_range(m, rank, world)
    idx, d1, d2 = knn2_fn(query[b:e], train)
    if world == 1:
        return idx, d1, d2
    per = (m + world - 1) // world                      # pad shards to one size for all_gather
    local = np.full((3, per), -1, np.int32)
    local[0, :e - b], local[1, :e - b], local[2, :e - b] = idx, d1, d2
    t = torch.from_numpy(local)
    if device is not None:
        t = t.to(device)
    out = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(out, t, group=group)
    res = np.empty((3, m), np.int32)
    for r in range(world):
        rb, re = shard_range(m, r, world)
        res[:, rb:re] = out[r].cpu().numpy()[:, :re - rb]
    return res[0], res[1], res[2]
