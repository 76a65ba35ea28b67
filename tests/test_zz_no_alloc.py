"""orbfe_knn2_device_noalloc — the allocation-free kNN-2 entry point (scratch in a static device array; no
counterpart in the reference) against the existing
orbfe_knn2_device through the C ABI: bit-exact (integer work, the tolerance of every kNN test here is zero), on two
calls in a row that share the static scratch, plus the size limit.  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest

from orb_slam2_annotate_b200 import synth

pytestmark = pytest.mark.gpu

SCRATCH_BYTES = 8 << 20     # ORBFE_KNN2_NOALLOC_SCRATCH_BYTES


def _existing(orbfe, dq, nq, dt, nt, th, ratio):
    import torch
    mt = orbfe.ORBmatcher(ratio)
    mt.set_knn_engine("mma")
    out = torch.full((3, nq), -9, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    mt.knn2_device(dq.data_ptr(), nq, dt.data_ptr(), nt, out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(), th, ratio)
    mt.sync()
    return out.cpu().numpy()


def _noalloc(orbfe, dq, nq, dt, nt, th, ratio):
    import torch
    L = orbfe.lib()
    out = torch.full((3, nq), -9, dtype=torch.int32, device="cuda")
    st = torch.cuda.Stream()
    torch.cuda.synchronize()
    rc = L.orbfe_knn2_device_noalloc(0, C.c_void_p(st.cuda_stream), dq.data_ptr(), nq, dt.data_ptr(), nt, th, ratio,
                                     out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr())
    assert rc == 0, L.orbfe_last_error()
    st.synchronize()
    return out.cpu().numpy()


def test_knn2_noalloc_matches_existing_entry_point_twice(orbfe, oracle):
    import torch
    torch.cuda.set_device(0)
    # call 1: 40 000 x 30 000 (one train segment); call 2: fewer queries against more rows (19 segments), so its
    # partial results land on top of what call 1 left in the shared static array
    for nq, nt, seed in ((40000, 30000, 1), (2000, 60000, 5)):
        t = synth.descriptors(nt, seed)
        q = synth.planted_descriptors(t, nq, seed + 1)
        dq, dt = torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda()
        want = _existing(orbfe, dq, nq, dt, nt, 50, 0.7)
        got = _noalloc(orbfe, dq, nq, dt, nt, 50, 0.7)
        assert np.array_equal(got, want), (nq, nt)
        assert (got[0] >= 0).sum() > 0
        if nq <= 2000:                                   # and both equal the CPU oracle
            oi, o1, o2 = oracle.knn2(q, t, 50, 0.7)
            assert np.array_equal(got[0], oi) and np.array_equal(got[1], o1) and np.array_equal(got[2], o2)


def test_knn2_noalloc_rejects_sizes_beyond_the_static_scratch(orbfe):
    import torch
    L = orbfe.lib()
    d = torch.zeros(64, dtype=torch.uint8, device="cuda")
    # 800 000 queries x 1 100 000 train rows: two train segments -> 19.2 MB of partial results > 8 MiB
    nq, nt = 800000, 1100000
    assert 3 * 4 * nq * 2 > SCRATCH_BYTES
    rc = L.orbfe_knn2_device_noalloc(0, None, d.data_ptr(), nq, d.data_ptr(), nt, 50, 0.7, d.data_ptr(), d.data_ptr(), d.data_ptr())
    assert rc == -3                                       # ORBFE_ERR_CAPACITY, nothing was launched
    assert b"scratch" in L.orbfe_last_error()
    assert L.orbfe_knn2_device_noalloc(0, None, None, 5, d.data_ptr(), 1, 50, 0.7, d.data_ptr(), d.data_ptr(), d.data_ptr()) == -1
    torch.cuda.synchronize()
