"""GPU parity of the entry points bench.py times (VERDICT r01 "test what the bench times"):
orbfe_search_init_consecutive_device (multi-pair launch, prev == NULL branch, scratch carve for nframes - 1 pairs),
orbfe_knn2_device on device buffers with both engines (integer-pipe popc kernel incl. its Q = 4 variant and the
> 2^20-row segment merge; tcgen05 tensor-core kernel incl. its train segments), and the query-sharded search of
BASELINE.json configs[4] through orbfe_knn2_sharded.  Everything bit-exact against the CPU oracle
(src/ORBmatcher.cc:469-600, :233-269, :1828-1844)."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from orb_slam2_annotate_b200 import synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _consecutive(orbfe, oracle, frames, nfeat=2000, window=100, ratio=0.9, ori=True):
    import torch
    B, H, W = frames.shape
    ex = orbfe.ORBextractor(nfeat, 1.2, 8, 20, 7)
    kps, desc, counts = ex.extract_batch(frames)
    kp_ptr, desc_ptr, cnt_ptr, cap = ex.device_results()
    mt = orbfe.ORBmatcher(ratio, ori)
    d_m12 = torch.full((max(B - 1, 1), cap), -7, dtype=torch.int32, device="cuda")
    d_nm = torch.full((max(B - 1, 1),), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    mt.search_init_consecutive_device(kp_ptr, desc_ptr, cnt_ptr, cap, B, W, H, d_m12.data_ptr(), d_nm.data_ptr(), window)
    mt.sync()
    m12, nm = d_m12.cpu().numpy(), d_nm.cpu().numpy()
    scale = ex.GetScaleFactors()
    total = 0
    for p in range(B - 1):
        n1, n2 = int(counts[p]), int(counts[p + 1])
        k1, d1, k2, d2 = kps[p][:n1], desc[p][:n1], kps[p + 1][:n2], desc[p + 1][:n2]
        prev = np.ascontiguousarray(np.stack([k1["x"], k1["y"]], 1), np.float32)      # prev == NULL: F1's own positions
        o1 = oracle.frame_view(k1, d1, W, H, scale); o2 = oracle.frame_view(k2, d2, W, H, scale)
        no, mo = oracle.search_for_initialization(o1, o2, prev, window, ratio, ori)
        assert int(nm[p]) == no, (p, int(nm[p]), no)
        assert np.array_equal(m12[p][:n1], mo), p
        total += no
    return total, counts


def test_search_init_consecutive_device_16_frames(orbfe, oracle):
    frames, _ = synth.sequence(16, 640, 480, seed=5)
    total, counts = _consecutive(orbfe, oracle, frames)
    assert total > 15 * 50 and counts.min() > 500


def test_search_init_consecutive_device_empty_frame_and_two_frames(orbfe, oracle):
    frames, _ = synth.sequence(5, 640, 480, seed=6)
    frames = frames.copy()
    frames[2] = 128                                    # constant image: no keypoints -> pairs (1,2) and (2,3) are empty
    total, counts = _consecutive(orbfe, oracle, frames)
    assert counts[2] == 0 and total > 0
    total2, _ = _consecutive(orbfe, oracle, np.ascontiguousarray(frames[:2]), nfeat=1000, window=40, ratio=0.7, ori=False)
    assert total2 > 0


def _knn_device(orbfe, q, t, engine, th=50, ratio=0.7):
    import torch
    mt = orbfe.ORBmatcher(ratio)
    mt.set_knn_engine(engine)
    dq = torch.from_numpy(q).cuda()
    dt = torch.from_numpy(t).cuda() if len(t) else torch.zeros((1, 32), dtype=torch.uint8, device="cuda")
    out = torch.full((3, len(q)), -9, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    mt.knn2_device(dq.data_ptr(), len(q), dt.data_ptr(), len(t), out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(), th, ratio)
    mt.sync()
    o = out.cpu().numpy()
    return o[0], o[1], o[2]


@pytest.mark.parametrize("engine", ["popc", "mma", "fp4"])
@pytest.mark.parametrize("m,n", [(1, 1), (7, 1), (300, 127), (257, 129), (1000, 20000), (5000, 777), (4096, 4096), (70000, 3000)])
def test_knn2_device_engines_match_oracle(orbfe, oracle, engine, m, n):
    t = synth.descriptors(n, 1)
    q = synth.planted_descriptors(t, m, 2)
    idx, d1, d2 = _knn_device(orbfe, q, t, engine)
    oi, o1, o2 = oracle.knn2(q, t, 50, 0.7)
    assert np.array_equal(d1, o1) and np.array_equal(d2, o2) and np.array_equal(idx, oi)


@pytest.mark.parametrize("engine", ["popc", "mma", "fp4"])
def test_knn2_device_ties_extremes_empty(orbfe, oracle, engine):
    t = np.zeros((300, 32), np.uint8)
    t[10:] = 0xFF
    q = np.zeros((5, 32), np.uint8)
    q[1] = 0xFF          # distance 256 to rows 0..9, 0 to rows 10..
    q[2, 0] = 1          # distance 1 to rows 0..9: lowest index wins, d2 == d1
    q[3, :16] = 0xFF     # distance 128 to everything
    idx, d1, d2 = _knn_device(orbfe, q, t, engine, th=50, ratio=0.9)
    oi, o1, o2 = oracle.knn2(q, t, 50, 0.9)
    assert np.array_equal(idx, oi) and np.array_equal(d1, o1) and np.array_equal(d2, o2)
    idx, d1, d2 = _knn_device(orbfe, q, np.zeros((0, 32), np.uint8), engine)
    assert (idx == -1).all() and (d1 == 256).all() and (d2 == 256).all()


@pytest.mark.parametrize("engine", ["popc", "mma", "fp4"])
def test_knn2_device_q4_variant(orbfe, oracle, engine):
    """nq >= 303 104 takes knn2_popc_kernel<4>; the tensor-core engine runs 1250 query tiles x 1 segment."""
    t = synth.descriptors(2000, 3)
    rng = np.random.default_rng(4)
    q = rng.integers(0, 256, (320000, 32), dtype=np.uint8)
    rows = rng.choice(len(q), 32000, replace=False)
    q[rows] = t[rng.integers(0, len(t), len(rows))] ^ (rng.random((len(rows), 32)) < 0.03).astype(np.uint8)
    idx, d1, d2 = _knn_device(orbfe, q, t, engine)
    oi, o1, o2 = oracle.knn2(q, t, 50, 0.7)
    assert np.array_equal(d1, o1) and np.array_equal(d2, o2) and np.array_equal(idx, oi)
    assert (idx >= 0).sum() > 10000


@pytest.mark.parametrize("engine", ["popc", "mma", "fp4"])
def test_knn2_device_train_over_2_20_rows_ties_across_segments(orbfe, oracle, engine):
    """1 100 000 train rows (> 2^20: the popc kernel merges two index segments, the tensor-core engine 148 train
    segments of 7 552 rows) x 64 queries whose best distance is reached by three rows planted in different segments:
    the lowest train index must win and d2 == d1."""
    n = 1_100_000
    rng = np.random.default_rng(9)
    t = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    q = rng.integers(0, 256, (64, 32), dtype=np.uint8)
    spots = [(1 << 20) - 1, 1 << 20, (1 << 20) + 1, 7551, 7552, 15104, n - 1, 0, 7487, 7488]     # 7488: the FP4 engine's segment length here
    for i in range(64):
        near = q[i].copy()
        near[i % 32] ^= np.uint8(1 << (i % 8))                       # distance 1
        a, b = spots[i % len(spots)], spots[(i + 3) % len(spots)]
        c = int(rng.integers(0, n))
        for pos in (a, b, c):                                        # later queries overwrite the fixed spots: the oracle decides
            t[pos] = near
    idx, d1, d2 = _knn_device(orbfe, q, t, engine)
    oi, o1, o2 = oracle.knn2(q, t, 50, 0.7)
    assert np.array_equal(d1, o1) and np.array_equal(d2, o2) and np.array_equal(idx, oi)
    assert (d1 <= 1).sum() >= 8 and (d2[d1 <= 1] <= 1).sum() >= 1    # ties were really planted


def test_knn2_sharded_world1(orbfe, oracle):
    """orbfe_knn2_sharded on one rank (a one-rank NCCL communicator): shard == all queries, packed 8-byte rows."""
    import torch
    t = synth.descriptors(3000, 1)
    q = synth.planted_descriptors(t, 5001, 2)
    comm = orbfe.Comm.from_torch(0)
    mt = orbfe.ORBmatcher(0.7)
    dq, dt = torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda()
    per = orbfe.shard_rows(len(q), 1)
    out = torch.zeros((per, 8), dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    mt.knn2_sharded(comm, dq.data_ptr(), len(q), dt.data_ptr(), len(t), out.data_ptr(), 50, 0.7)
    mt.sync()
    idx, d1, d2 = orbfe.unpack_results(out.cpu().numpy(), len(q))
    oi, o1, o2 = oracle.knn2(q, t, 50, 0.7)
    assert np.array_equal(idx, oi) and np.array_equal(d1, o1) and np.array_equal(d2, o2)
    comm.close()


def test_knn2_sharded_world2_nccl(orbfe):
    """Two ranks on two GPUs (skipped on a one-GPU box): tools/knn_sharded.py under torch.distributed.run checks
    sharded == single GPU == oracle subsample and the planted property, and exits non-zero otherwise."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", str(port), os.path.join(ROOT, "tools", "knn_sharded.py"), "--nq", "200000", "--nt", "150000",
                        "--check"], capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert '"parity": "ok"' in r.stdout
