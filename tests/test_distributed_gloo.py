"""world_size-2 gloo tests of the multi-GPU host logic (query sharding + result all_gather,
frame round-robin).  The per-rank compute is the CPU oracle here; on the GPU box the same
function is driven with ORBmatcher.knn2 (bench.py)."""
import os
import socket

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

from orb_slam2_annotate_b200 import synth
from orb_slam2_annotate_b200.distributed import frames_for_rank, knn2_sharded, shard_range


def test_shard_range_covers_everything():
    for n in (0, 1, 7, 1000, 1_000_003):
        for world in (1, 2, 3, 8):
            parts = [shard_range(n, r, world) for r in range(world)]
            assert parts[0][0] == 0 and parts[-1][1] == n
            assert all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
            sizes = [e - b for b, e in parts]
            assert max(sizes) - min(sizes) <= 1
    assert sorted(np.concatenate([frames_for_rank(10, r, 4) for r in range(4)]).tolist()) == list(range(10))


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle_py as O
    t = synth.descriptors(700, 1)
    qd = synth.planted_descriptors(t, 1001, 2)
    idx, d1, d2 = knn2_sharded(lambda a, b: O.knn2(a, b, 50, 0.7), qd, t)
    oi, o1, o2 = O.knn2(qd, t, 50, 0.7)
    ok = bool(np.array_equal(idx, oi) and np.array_equal(d1, o1) and np.array_equal(d2, o2))
    q.put((rank, ok, int((idx >= 0).sum())))
    dist.destroy_process_group()


def test_knn2_sharded_gloo_world2():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    res = [q.get(timeout=120) for _ in range(2)]
    [p.join(timeout=60) for p in procs]
    assert all(ok for _, ok, _ in res) and res[0][2] == res[1][2] > 0


# ---- the device path's wire format (orbfe_knn2_sharded, csrc/api_comm.cu): contiguous shards of shard_rows queries,
#      8-byte orbfe_knn_result rows, padding rows of 0xFF, one in-place all-gather

def test_shard_rows_matches_the_c_abi_and_rows_are_8_bytes():
    from orb_slam2_annotate_b200 import _lib
    from orb_slam2_annotate_b200.distributed import KNN_RESULT_DTYPE, shard_rows
    L = _lib.lib()                                   # loads without a GPU; orbfe_knn2_shard_rows is arithmetic only
    assert KNN_RESULT_DTYPE.itemsize == 8
    for n in (0, 1, 7, 1000, 1_000_000, 1_000_003):
        for world in (1, 2, 3, 8):
            assert L.orbfe_knn2_shard_rows(n, world) == shard_rows(n, world)
            assert shard_rows(n, world) * world >= n


def _wire_worker(rank, world, port, q):
    import torch
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle_py as O
    from orb_slam2_annotate_b200.distributed import KNN_RESULT_DTYPE, shard_rows, unpack_results
    t = synth.descriptors(900, 3)
    qd = synth.planted_descriptors(t, 1001, 4)           # 1001 queries over 2 ranks: shards of 501 and 500 + one padding row
    per = shard_rows(len(qd), world)
    b, e = rank * per, min((rank + 1) * per, len(qd))
    idx, d1, d2 = O.knn2(qd[b:e], t, 50, 0.7)             # stands in for the rank's orbfe_knn2_device
    buf = np.full(world * per, 0xFF, np.uint8).repeat(8).view(KNN_RESULT_DTYPE)   # what the library memsets
    buf["idx"][b:e], buf["d1"][b:e], buf["d2"][b:e] = idx, d1, d2
    mine = torch.from_numpy(buf.view(np.uint8).reshape(world, per * 8)[rank].copy())
    parts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine)                          # ncclAllGather of per * 8 bytes per rank on the GPU box
    gi, g1, g2 = unpack_results(torch.cat(parts).numpy(), len(qd))
    oi, o1, o2 = O.knn2(qd, t, 50, 0.7)
    q.put((rank, bool(np.array_equal(gi, oi) and np.array_equal(g1, o1) and np.array_equal(g2, o2)), int((gi >= 0).sum())))
    dist.destroy_process_group()


def test_device_wire_format_gloo_world2():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_wire_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    res = [q.get(timeout=120) for _ in range(2)]
    [p.join(timeout=60) for p in procs]
    assert all(ok for _, ok, _ in res) and res[0][2] == res[1][2] > 0
