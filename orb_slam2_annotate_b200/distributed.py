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


# ---- configs[4] on the device, behind the C ABI (orbfe_comm_*, orbfe_knn2_sharded) ---------------------------

KNN_RESULT_DTYPE = np.dtype([("idx", np.int32), ("d1", np.uint16), ("d2", np.uint16)])   # orbfe_knn_result, 8 bytes


def shard_rows(n: int, world: int) -> int:
    """Rows per shard of the device path (orbfe_knn2_shard_rows): rank r owns [r * per, min((r + 1) * per, n))."""
    return (n + world - 1) // world if n > 0 and world > 0 else 0


class Comm:
    """The library's NCCL communicator (``orbfe_comm``).  ``Comm.from_torch()`` builds one for the ranks of the
    default torch.distributed group: rank 0 draws the NCCL unique id through the C ABI and the 128 bytes travel by
    ``dist.broadcast`` (any backend; gloo or nccl), then every rank calls ncclCommInitRank inside the library."""

    def __init__(self, unique_id: bytes, world: int, rank: int, device: int):
        import ctypes as C
        from . import _lib
        self._L = _lib.lib()
        self._h = C.c_void_p()
        buf = (C.c_uint8 * 128).from_buffer_copy(unique_id)
        _lib.check(self._L.orbfe_comm_create(buf, world, rank, device, C.byref(self._h)), "orbfe_comm_create")
        self.world, self.rank, self.device = world, rank, device

    @staticmethod
    def unique_id() -> bytes:
        import ctypes as C
        from . import _lib
        buf = (C.c_uint8 * 128)()
        _lib.check(_lib.lib().orbfe_comm_unique_id(buf), "orbfe_comm_unique_id")
        return bytes(buf)

    @classmethod
    def from_torch(cls, device: int, group=None):
        import torch
        import torch.distributed as dist
        if not dist.is_initialized():
            return cls(cls.unique_id(), 1, 0, device)
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        on_gpu = dist.get_backend(group) == "nccl"
        t = torch.zeros(128, dtype=torch.uint8, device=f"cuda:{device}" if on_gpu else "cpu")
        if rank == 0:
            t.copy_(torch.frombuffer(bytearray(cls.unique_id()), dtype=torch.uint8))
        dist.broadcast(t, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        return cls(bytes(t.cpu().numpy().tobytes()), world, rank, device)

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.orbfe_comm_destroy(self._h)
            self._h.value = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def unpack_results(packed: np.ndarray, n: int):
    """Gathered orbfe_knn_result rows (padded to world * shard_rows) -> (idx, d1, d2) int32 of the n queries."""
    r = packed.view(KNN_RESULT_DTYPE).reshape(-1)[:n]
    return r["idx"].astype(np.int32), r["d1"].astype(np.int32), r["d2"].astype(np.int32)
