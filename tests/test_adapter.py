"""The C++ host side (include/orbfe_adapter.hpp: ORB_SLAM2::ORBextractor / ORBmatcher with the
reference's signatures) compiles against a cv:: type shim and, on the GPU box, produces the same
keypoints, descriptors and initialisation matches as the Python host side and the oracle."""
import os
import subprocess

import numpy as np
import pytest

from orb_slam2_annotate_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "cpp", "adapter_smoke")
EXE_GUIDED = os.path.join(ROOT, "tests", "cpp", "adapter_guided")


def _build(exe=EXE):
    subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "liborb_oracle.so"], check=True, stdout=subprocess.DEVNULL)
    subprocess.run(["g++", "-std=c++11", "-O2", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"),
                    "-I" + os.path.join(ROOT, "oracle", "cvshim"), "-I" + os.path.join(ROOT, "oracle"), "-o", exe, exe + ".cpp",
                    "-L" + os.path.join(ROOT, "orb_slam2_annotate_b200"), "-lorbfe", "-L" + os.path.join(ROOT, "oracle"), "-lorb_oracle",
                    "-Wl,-rpath," + os.path.join(ROOT, "orb_slam2_annotate_b200"), "-Wl,-rpath," + os.path.join(ROOT, "oracle")],
                   check=True)


@pytest.mark.parametrize("exe", [EXE, EXE_GUIDED])
def test_adapter_compiles_and_links(orbfe, exe):
    """adapter_guided instantiates every map-point dependent template of the header on stand-in types."""
    _build(exe)
    assert os.path.exists(exe)


@pytest.mark.gpu
def test_adapter_end_to_end(orbfe, oracle, tmp_path):
    _build()
    imgs, _ = synth.sequence(2, 640, 480, seed=4)
    inp, out = str(tmp_path / "in.bin"), str(tmp_path / "out.bin")
    with open(inp, "wb") as f:
        f.write(np.array([640, 480], np.int32).tobytes()); f.write(imgs.tobytes())
    r = subprocess.run([EXE, inp, out], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    print(r.stdout.strip())
    raw = open(out, "rb").read()
    off = 0
    frames = []
    for i in range(2):
        n = int(np.frombuffer(raw, np.int32, 1, off)[0]); off += 4
        k = np.frombuffer(raw, orbfe.KP_DTYPE, n, off); off += 28 * n
        d = np.frombuffer(raw, np.uint8, 32 * n, off).reshape(n, 32); off += 32 * n
        frames.append((k, d))
    nm, n1 = (int(v) for v in np.frombuffer(raw, np.int32, 2, off)); off += 8
    m12 = np.frombuffer(raw, np.int32, n1, off); off += 4 * n1
    t = oracle.OracleExtractor(2000).tables()
    ref = [oracle.OracleExtractor(2000)(im) for im in imgs]
    for (k, d), (ko, do) in zip(frames, ref):
        assert len(k) == len(ko)
        for f in ("x", "y", "size", "response", "octave", "class_id"):
            assert np.array_equal(k[f], ko[f])
        assert np.array_equal(k["angle"], ko["angle"]) and np.array_equal(d, do)
    prev = np.ascontiguousarray(np.stack([ref[0][0]["x"], ref[0][0]["y"]], 1), np.float32)
    no, mo = oracle.search_for_initialization(oracle.frame_view(ref[0][0], ref[0][1], 640, 480, t["scale"]),
                                              oracle.frame_view(ref[1][0], ref[1][1], 640, 480, t["scale"]), prev, 100, 0.9, True)
    assert nm == no and np.array_equal(m12, mo) and nm > 100
    # Frame epilogue through the adapter (TUM1 camera): bounds, mvKeysUn and mGrid of frame 0
    cam = np.array([517.306408, 516.469215, 318.643040, 255.313989], np.float32)
    dist = np.array([0.262383, -0.953104, -0.005358, 0.002628, 1.163314], np.float32)
    bnd = np.frombuffer(raw, np.float32, 4, off); off += 16
    assert bnd.tobytes() == oracle.image_bounds(cam, dist, 640, 480).tobytes()
    k0 = frames[0][0]
    un = np.frombuffer(raw, np.float32, 2 * len(k0), off).reshape(-1, 2); off += 8 * len(k0)
    want = oracle.undistort_points(np.stack([k0["x"], k0["y"]], 1), cam, dist)
    assert un.tobytes() == want.tobytes()
    ku = k0.copy(); ku["x"], ku["y"] = want[:, 0], want[:, 1]
    cs, items = oracle.assign_grid(oracle.frame_view(ku, frames[0][1], 640, 480, t["scale"], bounds=bnd))
    for c in range(64 * 48):
        cnt = int(np.frombuffer(raw, np.int32, 1, off)[0]); off += 4
        got = np.frombuffer(raw, np.int32, cnt, off); off += 4 * cnt
        assert np.array_equal(got, items[cs[c]:cs[c + 1]])
    # both SearchByBoW overloads through the adapter: counts against the oracle on the same stand-in feature vectors
    nb1, cnt1, len1, nb2, cnt2, len2 = (int(v) for v in np.frombuffer(raw, np.int32, 6, off)); off += 24
    assert off == len(raw)
    sides = []
    for k, d in frames:
        fv = {}
        for i in range(len(k)):
            fv.setdefault(int((d[i, 0] & 15) | (k["octave"][i] << 4)), []).append(i)
        idx = np.arange(len(k))
        valid = ((idx % 3 != 0) & ~((idx % 7 == 0) & (idx % 3 != 0))).astype(np.uint8)
        sides.append((d, k["angle"], fv, valid))
    (da, aa, fva, va), (db, ab, fvb, vb) = sides
    n1, _, o21 = oracle.search_by_bow(oracle.bow_view(da, aa, fva, va), oracle.bow_view(db, ab, fvb), 50, 0.75, True)
    n2, o12, _ = oracle.search_by_bow(oracle.bow_view(da, aa, fva, va), oracle.bow_view(db, ab, fvb, vb), 49, 0.75, True)
    assert (nb1, cnt1, len1) == (n1, int((o21 >= 0).sum()), len(db)) and n1 > 30
    assert (nb2, cnt2, len2) == (n2, int((o12 >= 0).sum()), len(da)) and n2 > 10


def _blobs(raw):
    out, off = [], 0
    while off < len(raw):
        n = int(np.frombuffer(raw, np.int32, 1, off)[0]); off += 4
        out.append(raw[off:off + n]); off += n
    return out


@pytest.mark.gpu
def test_adapter_guided_searches(orbfe, oracle, tmp_path):
    """tests/cpp/adapter_guided.cpp calls the four SearchByProjection overloads, both Fuse overloads, SearchBySim3 and
    SearchForTriangulation through the header on stand-in Frame / KeyFrame / MapPoint types (identity rotations, so the
    projections are redone here in float32); every result is checked against the oracle plus the reference's
    bookkeeping replayed on the oracle's search results."""
    _build(EXE_GUIDED)
    W, H = 640, 480
    imgs, offs = synth.sequence(2, W, H, seed=4)
    sx, sy = int(offs[0][0] - offs[1][0]), int(offs[0][1] - offs[1][1])        # a frame-0 pixel appears at x + sx in frame 1
    inp, out = str(tmp_path / "in.bin"), str(tmp_path / "out.bin")
    with open(inp, "wb") as f:
        f.write(np.array([W, H, sx, sy], np.int32).tobytes()); f.write(imgs.tobytes())
    r = subprocess.run([EXE_GUIDED, inp, out], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    print(r.stdout.strip())
    b = _blobs(open(out, "rb").read())
    assert len(b) == 34
    i32 = lambda x: np.frombuffer(x, np.int32)
    f32 = np.float32
    k0, k1 = (np.frombuffer(b[i], orbfe.KP_DTYPE) for i in (0, 2))
    d0, d1 = (np.frombuffer(b[i], np.uint8).reshape(-1, 32) for i in (1, 3))
    N0, N1 = len(k0), len(k1)
    rows = np.frombuffer(b[4], f32).reshape(-1, 8)
    pos, nz, level, nobs0, bad0 = rows[:, :3], rows[:, 3], rows[:, 4].astype(np.int32), rows[:, 5].astype(np.int32), rows[:, 6] > 0
    pdesc = np.frombuffer(b[5], np.uint8).reshape(-1, 32)
    A, B = i32(b[6]), i32(b[7])
    assert len(A) == N0 and len(B) == N1 and N0 > 500 and N1 > 500
    a_index = {int(v): i for i, v in enumerate(A) if v >= 0}
    b_index = {int(v): j for j, v in enumerate(B) if v >= 0}
    t = oracle.OracleExtractor(1000).tables()
    scale = t["scale"]
    t1 = np.array([sx, sy, 0], f32)
    fr1 = lambda: oracle.frame_view(k1, d1, W, H, scale)
    kf0 = lambda: oracle.frame_view(k0, d0, W, H, scale, keyframe=True)
    kf1 = lambda: oracle.frame_view(k1, d1, W, H, scale, keyframe=True)
    in_frame = lambda u, v: (u >= 0) & (u <= W) & (v >= 0) & (v <= H)                # Frame: float bounds, closed
    in_kf = lambda u, v: (u >= 0) & (u < W) & (v >= 0) & (v < H)                     # KeyFrame::IsInImage

    def expect(pre, res, ids):
        post = pre.copy()
        post[res >= 0] = ids[res[res >= 0]]
        post[res == -2] = -1
        return post

    # case A: SearchByProjection(Frame&, vector<MapPoint*>, th)
    local, pre, n, post = i32(b[8]), i32(b[9]), int(i32(b[10])[0]), i32(b[11])
    ai = np.array([a_index[int(v)] for v in local])
    px = (pos[local, 0] + f32(sx)).astype(f32); py = (pos[local, 1] + f32(sy)).astype(f32)
    mps = dict(track_in_view=((ai % 5 != 0) & ~bad0[local]).astype(np.uint8), proj_x=px, proj_y=py, proj_xr=(px - f32(4)).astype(f32),
               level=level[local], view_cos=np.where(ai % 2 == 1, f32(0.9995), f32(0.9)).astype(f32), desc=pdesc[local], observations=nobs0[local])
    no, res = oracle.search_by_projection(fr1(), mps, 3.0, 0.8, (pre >= 0).astype(np.uint8))
    assert n == no and np.array_equal(post, expect(pre, res, local)) and n > 100

    # case B: SearchByProjection(CurrentFrame, LastFrame, th, bMono)
    pre, n, post = i32(b[12]), int(i32(b[13])[0]), i32(b[14])
    idx = np.arange(N0)
    has = A >= 0
    Ai = np.where(has, A, 0)
    u = (pos[Ai, 0] + f32(sx)).astype(f32); v = (pos[Ai, 1] + f32(sy)).astype(f32)
    pts = dict(valid=(has & (idx % 11 != 0) & in_frame(u, v)).astype(np.uint8), u=u, v=v, ur=(u - f32(4)).astype(f32),
               radius=(f32(7) * scale[k0["octave"]]).astype(f32), min_level=k0["octave"] - 1, max_level=k0["octave"] + 1,
               angle=k0["angle"], desc=pdesc[Ai], observations=nobs0[Ai])
    no, res = oracle.search_by_projection_frame(fr1(), pts, True, (pre >= 0).astype(np.uint8))
    assert n == no and np.array_equal(post, expect(pre, res, A)) and n > 100

    # case C: SearchByProjection(Frame&, KeyFrame*, set<MapPoint*>, th, ORBdist)
    pre, n, post = i32(b[15]), int(i32(b[16])[0]), i32(b[17])
    pts = dict(valid=(has & ~bad0[Ai] & (idx % 7 != 0) & in_frame(u, v)).astype(np.uint8), u=u, v=v,
               radius=(f32(10) * scale[level[Ai]]).astype(f32), level=level[Ai], angle=k0["angle"], desc=pdesc[Ai])
    no, res = oracle.search_by_projection_keyframe(fr1(), pts, 100, True, (pre >= 0).astype(np.uint8))
    assert n == no and np.array_equal(post, expect(pre, res, A)) and n > 100

    # case D: SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th)
    loop, pre, n, post = i32(b[18]), i32(b[19]), int(i32(b[20])[0]), i32(b[21])
    lu = (pos[loop, 0] + f32(sx)).astype(f32); lv = (pos[loop, 1] + f32(sy)).astype(f32)
    seen = ~bad0[loop] & in_kf(lu, lv) & (nz[loop] > 0)
    pts = dict(valid=(seen & ~np.isin(loop, pre[pre >= 0])).astype(np.uint8), u=lu, v=lv, radius=(f32(10) * scale[level[loop]]).astype(f32),
               level=level[loop], desc=pdesc[loop])
    no, res = oracle.search_by_projection_sim3(kf1(), pts, (pre >= 0).astype(np.uint8))
    assert n == no and np.array_equal(post, expect(pre, res, loop)) and n > 50

    # case G: SearchBySim3
    pre, n, post = i32(b[22]), int(i32(b[23])[0]), i32(b[24])
    done1 = pre >= 0
    done2 = np.zeros(N1, bool)
    done2[[b_index[int(v)] for v in pre[done1]]] = True
    hasB = B >= 0
    Bi = np.where(hasB, B, 0)
    p1 = dict(valid=(has & ~done1 & ~bad0[Ai] & in_kf(u, v)).astype(np.uint8), u=u, v=v, radius=(f32(7.5) * scale[level[Ai]]).astype(f32),
              level=level[Ai], desc=pdesc[Ai])
    c2 = (pos[Bi] + t1).astype(f32)                                              # R2w * P + t2w
    c1 = (c2 + np.array([-sx, -sy, 0], f32)).astype(f32)                         # sR12 * p3Dc2 + t12
    p2 = dict(valid=(hasB & ~done2 & ~bad0[Bi] & in_kf(c1[:, 0], c1[:, 1])).astype(np.uint8), u=c1[:, 0].copy(), v=c1[:, 1].copy(),
              radius=(f32(7.5) * scale[level[Bi]]).astype(f32), level=level[Bi], desc=pdesc[Bi])
    no, res = oracle.search_by_sim3(kf0(), kf1(), p1, p2)
    assert n == no and np.array_equal(post, expect(pre, res, B)) and n > 50

    # case H: SearchForTriangulation
    n, pairs = int(i32(b[25])[0]), i32(b[26]).reshape(-1, 2)
    sides = []
    for k, d, hm in ((k0, d0, has), (k1, d1, hasB)):
        fv = {}
        for i in range(len(k)):
            fv.setdefault(int((d[i, 0] & 7) | (k["octave"][i] << 3)), []).append(i)
        sides.append(oracle.tri_side(k, d, hm.astype(np.uint8), fv))
    F12 = np.array([[0, 0, sy], [0, 0, -sx], [-sy, sx, 0]], f32)
    no, res = oracle.search_for_triangulation(sides[0], sides[1], F12, 300.0, 200.0, scale, t["sigma2"], False, False)
    assert n == no and np.array_equal(pairs, np.stack([np.nonzero(res >= 0)[0], res[res >= 0]], 1)) and n > 20

    # case F: Fuse(KeyFrame*, Scw, vpPoints, th, vpReplacePoint): the search, then :1208-1222 replayed
    n, replace, mps_post = int(i32(b[27])[0]), i32(b[28]), i32(b[29])
    pts = dict(valid=seen.astype(np.uint8), u=lu, v=lv, radius=(f32(4) * scale[level[loop]]).astype(f32), level=level[loop], desc=pdesc[loop])
    best, _ = oracle.fuse_search(kf1(), pts, None)
    mps_now, rep, nf = B.copy(), np.full(len(loop), -1, np.int32), 0
    for i in range(len(loop)):
        if best[i] < 0:
            continue
        cur = mps_now[best[i]]
        if cur >= 0:
            if not bad0[cur]:
                rep[i] = cur
        else:
            mps_now[best[i]] = loop[i]
        nf += 1
    assert n == nf and np.array_equal(replace, rep) and np.array_equal(mps_post, mps_now) and nf > 100
    assert (rep >= 0).sum() > 20 and (mps_now != B).sum() > 20

    # case E: Fuse(KeyFrame*, vector<MapPoint*>, th): the search with the chi-square gate, then :1057-1073 replayed
    cand, n, mps_post, state = i32(b[30]), int(i32(b[31])[0]), i32(b[32]), i32(b[33]).reshape(-1, 3)
    ci = np.where(cand >= 0, cand, 0)
    cu = (pos[ci, 0] + f32(sx)).astype(f32); cv_ = (pos[ci, 1] + f32(sy)).astype(f32)
    pts = dict(valid=((cand >= 0) & ~bad0[ci] & in_kf(cu, cv_) & (nz[ci] > 0)).astype(np.uint8), u=cu, v=cv_, ur=cu,
               radius=(f32(3) * scale[level[ci]]).astype(f32), level=level[ci], desc=pdesc[ci])
    best, _ = oracle.fuse_search(kf1(), pts, t["inv_sigma2"])
    bad, nobs, mps_now, nf = bad0.copy(), nobs0.copy(), B.copy(), 0
    kfidx = np.full(len(rows), -1, np.int32)
    kfidx[B[hasB]] = np.nonzero(hasB)[0]
    in_kf1 = np.zeros(len(rows), bool)
    in_kf1[B[hasB]] = True
    swaps = 0
    for c in range(len(cand)):
        p = cand[c]
        if best[c] < 0 or bad[p] or in_kf1[p]:
            continue
        cur = mps_now[best[c]]
        if cur >= 0:
            if not bad[cur]:
                if nobs[cur] > nobs[p]:
                    bad[p] = True                                  # pMP->Replace(pMPinKF): pMP is in no key frame here
                else:
                    bad[cur] = True                                # pMPinKF->Replace(pMP)
                    mps_now[best[c]] = p; in_kf1[p] = True; kfidx[p] = best[c]; nobs[p] += 1
                    swaps += 1
        else:
            mps_now[best[c]] = p; in_kf1[p] = True; kfidx[p] = best[c]; nobs[p] += 1
        nf += 1
    assert n == nf and np.array_equal(mps_post, mps_now) and nf > 100 and swaps > 5
    assert np.array_equal(state[:, 0] != 0, bad) and np.array_equal(state[:, 1], nobs) and np.array_equal(state[:, 2], kfidx)
