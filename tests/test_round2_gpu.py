"""GPU regression tests for the round-1 advisor findings and the round-2 entry points:
two live extractor handles of different feature budgets (per-function shared-memory attribute), the zero-copy level-0
alias with a batch below the handle's capacity and nothing mapped behind the caller's buffer, orbfe_pyramid_host,
orbfe_matcher_fetch, and the single-frame schedule (per-level fork / join, pinned staging) against the batched one."""
import ctypes as C

import numpy as np
import pytest

from orb_slam2_annotate_b200 import synth

pytestmark = pytest.mark.gpu


def _same(a, b):
    (ka, da), (kb, db) = a, b
    return len(ka) == len(kb) and ka.tobytes() == kb.tobytes() and np.array_equal(da, db)


def test_two_live_extractors_of_different_budgets_alternate(orbfe, oracle):
    """Tracking keeps ORBextractor(2 * nFeatures) and ORBextractor(nFeatures) alive (src/Tracking.cc:122-128): the
    quadtree kernel's dynamic shared-memory limit is per function and device, so the smaller handle must not lower it
    under the larger one."""
    img = synth.frame(640, 480, 3)
    big, small = orbfe.ORBextractor(2000), orbfe.ORBextractor(1000)
    ref_big, ref_small = oracle.OracleExtractor(2000)(img), oracle.OracleExtractor(1000)(img)
    for _ in range(3):
        kb, db = big(img)
        ks, ds = small(img)
        for (k, d), (ko, do) in (((kb, db), ref_big), ((ks, ds), ref_small)):
            assert len(k) == len(ko)
            assert k.tobytes() == ko.tobytes() and np.array_equal(d, do)
    # batched path of the large handle after the small one ran (eager launches, not the graph)
    imgs = np.ascontiguousarray(np.repeat(img[None], 6, 0))
    small.extract_batch(imgs)
    kps, desc, counts = big.extract_batch(imgs)
    assert all(int(c) == len(ref_big[0]) for c in counts)


def test_level0_alias_with_batch_below_capacity_and_tight_buffer(orbfe):
    """orbfe_extract_batch_device uses the caller's frames in place as level 0.  The handle was sized for 16 frames, the
    call brings 3 in a buffer with NOTHING behind the last row (the allocation ends exactly there, 720 rows so that the
    bottom FAST cells' 40-row TMA boxes reach past the frame): results must equal the copied path's."""
    import torch
    W, H, B = 1280, 720, 3
    imgs = np.stack([synth.frame(W, H, s) for s in range(B)])
    ex = orbfe.ORBextractor(1500, max_width=W, max_height=H, max_batch=16)
    ref = orbfe.ORBextractor(1500)
    want = [ref(imgs[b]) for b in range(B)]
    # an exact-size device allocation: torch's caching allocator rounds to 512 B, W * H * B is a multiple of it
    assert (W * H * B) % 512 == 0
    dev = torch.empty(W * H * B, dtype=torch.uint8, device="cuda")
    dev.copy_(torch.from_numpy(imgs.reshape(-1)))
    torch.cuda.synchronize()
    ex.extract_batch_device(dev.data_ptr(), B, W, H)
    ex.sync()
    kps, desc, counts = ex.fetch(B)
    for b in range(B):
        n = int(counts[b])
        assert _same((kps[b][:n], desc[b][:n]), want[b]), b
    # the same handle again with the full capacity (tensor maps re-encoded for 16 frames) and back to 3
    big = torch.from_numpy(np.ascontiguousarray(np.tile(imgs, (6, 1, 1))[:16])).cuda()
    ex.extract_batch_device(big.data_ptr(), 16, W, H); ex.sync()
    ex.extract_batch_device(dev.data_ptr(), B, W, H); ex.sync()
    kps, desc, counts = ex.fetch(B)
    assert _same((kps[1][:int(counts[1])], desc[1][:int(counts[1])]), want[1])


def test_pyramid_host_equals_per_level_download(orbfe):
    img = synth.frame(752, 480, 5)
    ex = orbfe.ORBextractor(1200)
    ex(img)
    L = orbfe.lib()
    n = ex.GetLevels()
    lv = (C.c_void_p * 16)(); w = (C.c_int32 * 16)(); h = (C.c_int32 * 16)(); st = (C.c_ssize_t * 16)()
    assert L.orbfe_pyramid_host(ex._h, 0, lv, w, h, st) == 0, L.orbfe_last_error()
    for l in range(n):
        want = ex.pyramid_level(l)
        buf = (C.c_uint8 * (st[l] * h[l])).from_address(lv[l])
        got = np.frombuffer(buf, np.uint8).reshape(h[l], st[l])[:, :w[l]]
        assert got.shape == want.shape and np.array_equal(got, want), l
    assert L.orbfe_pyramid_host(ex._h, 5, lv, w, h, st) == -1          # frame outside the last batch


def test_matcher_fetch_is_ordered_behind_the_search(orbfe, oracle):
    import torch
    t = synth.descriptors(5000, 1)
    q = synth.planted_descriptors(t, 3000, 2)
    mt = orbfe.ORBmatcher(0.7)
    dq, dt = torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda()
    out = torch.full((3, len(q)), -5, dtype=torch.int32, device="cuda")
    host = orbfe.pinned_empty((3, len(q)), np.int32)
    host[...] = -9
    torch.cuda.synchronize()
    mt.knn2_device(dq.data_ptr(), len(q), dt.data_ptr(), len(t), out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(), 50, 0.7)
    mt.fetch(host.ctypes.data, out.data_ptr(), host.nbytes)
    mt.sync()
    oi, o1, o2 = oracle.knn2(q, t, 50, 0.7)
    assert np.array_equal(host[0], oi) and np.array_equal(host[1], o1) and np.array_equal(host[2], o2)
    assert orbfe.lib().orbfe_matcher_fetch(mt._h, None, out.data_ptr(), 16) == -1


def test_single_frame_schedule_equals_batched(orbfe):
    """Batches <= 4 replay a forked CUDA graph (FAST / quadtree / blur per level beside the pyramid chain) and stage
    pageable buffers through pinned memory; batches > 4 run the plain schedule.  Same frames, same bytes — also on the
    third call, which is the first graph replay, and through a row-strided view."""
    imgs, _ = synth.sequence(8, 640, 480, seed=11)
    ex1, ex8 = orbfe.ORBextractor(1000), orbfe.ORBextractor(1000)
    kps, desc, counts = ex8.extract_batch(imgs)
    for rep in range(3):
        for b in (0, 5):
            k, d = ex1(imgs[b])
            assert _same((k, d), (kps[b][:int(counts[b])], desc[b][:int(counts[b])])), (rep, b)
    wide = np.zeros((480, 704), np.uint8)
    wide[:, :640] = imgs[2]
    k, d = ex1(wide[:, :640])                                              # stride 704, pageable
    assert _same((k, d), (kps[2][:int(counts[2])], desc[2][:int(counts[2])]))
    k2, d2, c2 = ex1.extract_batch(imgs[:3])                               # batch 3: small path with three frames
    for b in range(3):
        assert _same((k2[b][:int(c2[b])], d2[b][:int(c2[b])]), (kps[b][:int(counts[b])], desc[b][:int(counts[b])]))
