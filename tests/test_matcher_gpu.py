"""GPU parity of the guided searches (C ABI) against the CPU oracle: SearchForInitialization,
SearchByProjection(Frame, MapPoints) and ComputeStereoMatches.  Match indices, counts and the
float outputs (updated vbPrevMatched, mvuRight, mvDepth) must be bit-exact."""
import numpy as np
import pytest

from orb_slam2_annotate_b200 import synth

pytestmark = pytest.mark.gpu

_cache = {}


def _frames(oracle, kind, seed, nf):
    key = (kind, seed, nf)
    if key not in _cache:
        if kind == "mono":
            imgs, _ = synth.sequence(3, 640, 480, seed=seed)
        else:
            imgs = np.stack(synth.stereo_pair(1241, 376, seed=seed, disparity=20))
        out = []
        for im in imgs:
            oc = oracle.OracleExtractor(nf)
            k, d = oc(im)
            out.append((k, d, [oc.level(l) for l in range(8)], oc.tables()))
        _cache[key] = out
    return _cache[key]


@pytest.mark.parametrize("seed,ratio,ori,window", [(0, 0.9, True, 100), (1, 0.9, False, 100), (2, 0.7, True, 40), (3, 0.9, True, 300)])
def test_search_for_initialization(orbfe, oracle, seed, ratio, ori, window):
    fr = _frames(oracle, "mono", seed, 2000)
    (k1, d1, _, t), (k2, d2, _, _), (k3, d3, _, _) = fr
    F1 = orbfe.Frame(k1, d1, 640, 480, t["scale"]); F2 = orbfe.Frame(k2, d2, 640, 480, t["scale"])
    F3 = orbfe.Frame(k3, d3, 640, 480, t["scale"])
    mt = orbfe.ORBmatcher(ratio, ori)
    prev_g = np.ascontiguousarray(np.stack([k1["x"], k1["y"]], 1), np.float32)
    prev_o = prev_g.copy()
    for Fg, (ko, do) in ((F2, (k2, d2)), (F3, (k3, d3))):     # second call reuses the updated vbPrevMatched
        ng, mg = mt.SearchForInitialization(F1, Fg, prev_g, window)
        o1 = oracle.frame_view(k1, d1, 640, 480, t["scale"]); o2 = oracle.frame_view(ko, do, 640, 480, t["scale"])
        no, mo = oracle.search_for_initialization(o1, o2, prev_o, window, ratio, ori)
        assert ng == no and np.array_equal(mg, mo) and prev_g.tobytes() == prev_o.tobytes()
    assert ng > 50


def test_search_init_degenerate(orbfe, oracle):
    fr = _frames(oracle, "mono", 0, 2000)
    (k1, d1, _, t), (k2, d2, _, _), _ = fr
    mt = orbfe.ORBmatcher(0.9, True)
    empty = orbfe.Frame(k1[:0], d1[:0], 640, 480, t["scale"])
    F1 = orbfe.Frame(k1, d1, 640, 480, t["scale"])
    prev = np.ascontiguousarray(np.stack([k1["x"], k1["y"]], 1), np.float32)
    n, m = mt.SearchForInitialization(F1, empty, prev, 100)
    assert n == 0 and (m == -1).all()
    n, m = mt.SearchForInitialization(empty, F1, np.zeros((0, 2), np.float32), 100)
    assert n == 0 and len(m) == 0
    # identical frames: every octave-0 keypoint matches itself at distance 0
    n, m = mt.SearchForInitialization(F1, F1, prev.copy(), 100)
    o1 = oracle.frame_view(k1, d1, 640, 480, t["scale"])
    no, mo = oracle.search_for_initialization(o1, o1, prev.copy(), 100, 0.9, True)
    assert n == no and np.array_equal(m, mo)


def _mappoints(k, d, n, seed, scale):
    rng = np.random.default_rng(seed)
    src = rng.integers(0, len(k), n)
    desc = d[src].copy()
    flips = rng.integers(0, 41, n)
    for i in range(n):
        bits = rng.choice(256, flips[i], replace=False)
        np.bitwise_xor.at(desc[i], bits >> 3, (1 << (bits & 7)).astype(np.uint8))
    return dict(track_in_view=(rng.random(n) < 0.9).astype(np.uint8),
                proj_x=(k["x"][src] + rng.uniform(-3, 3, n)).astype(np.float32),
                proj_y=(k["y"][src] + rng.uniform(-3, 3, n)).astype(np.float32),
                proj_xr=(k["x"][src] - 20 + rng.uniform(-3, 3, n)).astype(np.float32),
                level=np.minimum(k["octave"][src] + rng.integers(0, 2, n), 7).astype(np.int32),
                view_cos=rng.choice(np.array([0.9, 0.999, 0.998], np.float32), n),
                desc=desc, observations=rng.integers(0, 3, n).astype(np.int32))


@pytest.mark.parametrize("nmp,th,stereo", [(20000, 1.0, False), (20000, 3.0, False), (5000, 5.0, True), (300, 15.0, False)])
def test_search_by_projection_local_points(orbfe, oracle, nmp, th, stereo):
    (k, d, _, t) = _frames(oracle, "mono", 5, 1000)[0]
    rng = np.random.default_rng(9)
    ur = np.where(rng.random(len(k)) < 0.5, k["x"] - 20, -1).astype(np.float32) if stereo else None
    mps = _mappoints(k, d, nmp, 11, t["scale"])
    occ_g = (rng.random(len(k)) < 0.05).astype(np.uint8); occ_o = occ_g.copy()
    F = orbfe.Frame(k, d, 640, 480, t["scale"], uright=ur)
    ng, mg = orbfe.ORBmatcher(0.8).SearchByProjection(F, mps, th, occ_g)
    no, mo = oracle.search_by_projection(oracle.frame_view(k, d, 640, 480, t["scale"], ur), mps, th, 0.8, occ_o)
    assert ng == no and np.array_equal(mg, mo) and np.array_equal(occ_g, occ_o)
    assert ng > 100


def _last_frame_points(orbfe, k_last, d_last, k_cur, seed, mode, th, stereo):
    """Last-frame map points placed so that the motion model projects them near their own
    keypoints: world point = back-projection of the last keypoint at a random depth; the current
    pose is the last one moved along z (forward / backward) or sideways (mono)."""
    rng = np.random.default_rng(seed)
    n = len(k_last)
    fx = fy = 520.0; cx, cy = 320.0, 240.0
    mbf = 40.0; mb = mbf / fx
    z = rng.uniform(4.0, 12.0, n).astype(np.float32)
    X = np.stack([(k_last["x"] - cx) / fx * z, (k_last["y"] - cy) / fy * z, z], 1).astype(np.float32)
    T_last = np.eye(4, dtype=np.float32)
    T_cur = np.eye(4, dtype=np.float32)
    T_cur[:3, 3] = {"forward": (0.01, 0.0, -0.3), "backward": (0.0, 0.01, 0.3), "mono": (0.05, 0.02, -0.3)}[mode]
    scale = np.float32(1.2) ** np.arange(8, dtype=np.float32)
    pts = orbfe.project_last_frame(T_cur, T_last, X, k_last["octave"], rng.random(n) < 0.9, rng.random(n) < 0.05,
                                   fx=fx, fy=fy, cx=cx, cy=cy, mbf=mbf, mb=mb, bounds=(0, 640, 0, 480),
                                   scale_factors=scale, th=th, mono=(mode == "mono"))
    assert pts["forward"] == (mode == "forward") and pts["backward"] == (mode == "backward")
    desc = d_last.copy()
    flips = rng.integers(0, 60, n)
    for i in range(n):
        bits = rng.choice(256, flips[i], replace=False)
        np.bitwise_xor.at(desc[i], bits >> 3, (1 << (bits & 7)).astype(np.uint8))
    pts.update(angle=k_last["angle"].astype(np.float32), desc=desc,
               observations=rng.integers(0, 3, n).astype(np.int32))
    return pts


@pytest.mark.parametrize("mode,th,stereo,ori", [("mono", 15.0, False, True), ("mono", 7.0, False, False),
                                                ("forward", 15.0, True, True), ("backward", 30.0, True, True)])
def test_search_by_projection_last_frame(orbfe, oracle, mode, th, stereo, ori):
    fr = _frames(oracle, "mono", 7, 2000)
    (kl, dl, _, t), (kc, dc, _, _) = fr[0], fr[1]
    rng = np.random.default_rng(3)
    ur = np.where(rng.random(len(kc)) < 0.5, kc["x"] - 5, -1).astype(np.float32) if stereo else None
    pts = _last_frame_points(orbfe, kl, dl, kc, 21, mode, th, stereo)
    occ = (rng.random(len(kc)) < 0.05).astype(np.uint8)
    F = orbfe.Frame(kc, dc, 640, 480, t["scale"], uright=ur)
    ng, mg = orbfe.ORBmatcher(0.9, ori).SearchByProjectionLast(F, pts, occ.copy())
    no, mo = oracle.search_by_projection_frame(oracle.frame_view(kc, dc, 640, 480, t["scale"], ur), pts, ori, occ)
    assert ng == no and np.array_equal(mg, mo)
    assert ng > 100 and (mo >= 0).sum() > 100
    if ori:
        assert (mo == -2).sum() > 0          # the rotation check removed something


def test_compute_stereo_matches(orbfe, oracle):
    for seed in (0, 1):
        (kl, dl, pl, t), (kr, dr, pr, _) = _frames(oracle, "stereo", seed, 2000)
        mbf = 386.1448; mb = mbf / 718.856
        L = orbfe.Frame(kl, dl, 1241, 376, t["scale"]); R = orbfe.Frame(kr, dr, 1241, 376, t["scale"])
        ug, dg = orbfe.ORBmatcher().ComputeStereoMatches(L, R, pl, pr, t["inv_scale"], mbf, mb)
        uo, do = oracle.compute_stereo_matches(oracle.frame_view(kl, dl, 1241, 376, t["scale"]),
                                               oracle.frame_view(kr, dr, 1241, 376, t["scale"]),
                                               oracle.pyramid_view(pl), oracle.pyramid_view(pr), t["inv_scale"], mbf, mb)
        assert ug.tobytes() == uo.tobytes() and dg.tobytes() == do.tobytes()
        assert (ug >= 0).sum() > 100


def test_compute_stereo_matches_batch_device(orbfe, oracle):
    """Three stereo pairs through the extractor as one interleaved batch, then ComputeStereoMatches on the
    device-resident keypoints / descriptors / pyramids: mvuRight and mvDepth bit-exact against the oracle."""
    pairs = [synth.stereo_pair(1241, 376, seed=s, disparity=20) for s in (0, 1, 2)]
    imgs = np.ascontiguousarray(np.stack([im for pr in pairs for im in pr]))
    ex = orbfe.ORBextractor(2000)
    kps, desc, counts = ex.extract_batch(imgs)
    mbf = 386.1448; mb = mbf / 718.856
    ug, dg = orbfe.ORBmatcher().ComputeStereoMatchesBatch(ex, 3, mbf, mb)
    for p, seed in enumerate((0, 1, 2)):
        if seed < 2:
            (kl, dl, pl, t), (kr, dr, pr, _) = _frames(oracle, "stereo", seed, 2000)
        else:
            fr = []
            for im in pairs[p]:
                oc = oracle.OracleExtractor(2000)
                k, d = oc(im)
                fr.append((k, d, [oc.level(l) for l in range(8)], oc.tables()))
            (kl, dl, pl, t), (kr, dr, pr, _) = fr
        nl = counts[2 * p]
        assert nl == len(kl) and counts[2 * p + 1] == len(kr)
        uo, do = oracle.compute_stereo_matches(oracle.frame_view(kl, dl, 1241, 376, t["scale"]),
                                               oracle.frame_view(kr, dr, 1241, 376, t["scale"]),
                                               oracle.pyramid_view(pl), oracle.pyramid_view(pr), t["inv_scale"], mbf, mb)
        assert ug[p, :nl].tobytes() == uo.tobytes() and dg[p, :nl].tobytes() == do.tobytes()
        assert (ug[p, nl:] == -1).all() and (dg[p, nl:] == -1).all()
        assert (ug[p, :nl] >= 0).sum() > 100


def _feat_vec(desc, octave, nbits, rng=None, drop=0.0):
    """Stand-in for DBoW2's FeatureVector: node id from the octave and the low `nbits` bits of the first
    descriptor byte, so that look-alike features of two frames mostly share a node."""
    node = (desc[:, 0].astype(np.int64) & ((1 << nbits) - 1)) | (octave.astype(np.int64) << nbits)
    fv = {}
    for i, k in enumerate(node):
        fv.setdefault(int(k) * 3 + 1, []).append(i)
    if rng is not None:
        for k in list(fv):
            if rng.random() < drop:
                del fv[k]                   # nodes present on one side only
            else:
                rng.shuffle(fv[k])          # vector order inside a node is arbitrary for the routine
    return fv


@pytest.mark.parametrize("pair,nbits,ratio,ori,shuffle", [(False, 4, 0.7, True, False), (True, 4, 0.75, True, True),
                                                           (False, 0, 0.9, False, True), (True, 1, 0.6, True, False),
                                                           (False, 6, 0.75, True, True)])
def test_search_by_bow(orbfe, oracle, pair, nbits, ratio, ori, shuffle):
    fr = _frames(oracle, "mono", 7, 2000)
    (k1, d1, _, _), (k2, d2, _, _) = fr[0], fr[1]
    rng = np.random.default_rng(31)
    fv1 = _feat_vec(d1, k1["octave"], nbits, rng if shuffle else None, 0.1)
    fv2 = _feat_vec(d2, k2["octave"], nbits, rng if shuffle else None, 0.1)
    v1 = (rng.random(len(k1)) < 0.8).astype(np.uint8)
    v2 = (rng.random(len(k2)) < 0.8).astype(np.uint8) if pair else None
    mt = orbfe.ORBmatcher(ratio, ori)
    ng, g12, g21 = mt.SearchByBoW(orbfe.BowSide(d1, k1["angle"], fv1, v1), orbfe.BowSide(d2, k2["angle"], fv2, v2), keyframe_pair=pair)
    no, o12, o21 = oracle.search_by_bow(oracle.bow_view(d1, k1["angle"], fv1, v1), oracle.bow_view(d2, k2["angle"], fv2, v2),
                                        49 if pair else 50, ratio, ori)
    assert ng == no and np.array_equal(g12, o12) and np.array_equal(g21, o21)
    assert ng == (o12 >= 0).sum() == (o21 >= 0).sum()
    assert ng > (5 if nbits == 6 else 40)


def test_search_by_bow_degenerate(orbfe, oracle):
    fr = _frames(oracle, "mono", 7, 2000)
    (k1, d1, _, _), (k2, d2, _, _) = fr[0], fr[1]
    mt = orbfe.ORBmatcher(0.7, True)
    fv1 = _feat_vec(d1, k1["octave"], 4)
    s1 = orbfe.BowSide(d1, k1["angle"], fv1)
    # no common node
    fv2 = {k + 1: v for k, v in _feat_vec(d2, k2["octave"], 4).items()}
    n, m12, m21 = mt.SearchByBoW(s1, orbfe.BowSide(d2, k2["angle"], fv2))
    assert n == 0 and (m12 == -1).all() and (m21 == -1).all()
    # an empty side / an empty feature vector
    n, m12, m21 = mt.SearchByBoW(s1, orbfe.BowSide(d2[:0], k2["angle"][:0], {}))
    assert n == 0 and len(m21) == 0 and (m12 == -1).all()
    n, m12, m21 = mt.SearchByBoW(orbfe.BowSide(d1, k1["angle"], {}), orbfe.BowSide(d2, k2["angle"], fv2))
    assert n == 0
    # a frame against itself: every feature finds itself at distance 0, second-best decides
    n, m12, m21 = mt.SearchByBoW(s1, s1)
    no, o12, o21 = oracle.search_by_bow(oracle.bow_view(d1, k1["angle"], fv1), oracle.bow_view(d1, k1["angle"], fv1), 50, 0.7, True)
    assert n == no and np.array_equal(m12, o12) and np.array_equal(m21, o21) and n > 500
    # malformed feature vectors are rejected
    bad = orbfe.BowSide(d2, k2["angle"], fv2)
    bad.indices[0] = len(d2)
    with pytest.raises(orbfe.OrbfeError):
        mt.SearchByBoW(s1, bad)


def test_search_by_bow_batch(orbfe, oracle):
    """Relocalisation shape: one frame against several key frames in one call, plus key-frame pairs."""
    fr = _frames(oracle, "mono", 7, 2000) + _frames(oracle, "mono", 5, 1000)
    rng = np.random.default_rng(77)
    sides, osides = [], []
    for (k, d, _, _) in fr:
        fv = _feat_vec(d, k["octave"], 4, rng, 0.05)
        valid = (rng.random(len(k)) < 0.85).astype(np.uint8)
        sides.append((orbfe.BowSide(d, k["angle"], fv, valid), orbfe.BowSide(d, k["angle"], fv)))
        osides.append((oracle.bow_view(d, k["angle"], fv, valid), oracle.bow_view(d, k["angle"], fv)))
    mt = orbfe.ORBmatcher(0.75, True)
    F = 1                                                   # the "current frame": same object in every pair
    kfs = [i for i in range(len(fr)) if i != F]
    res = mt.SearchByBoWBatch([sides[i][0] for i in kfs], [sides[F][1]] * len(kfs))
    assert len(res) == len(kfs)
    for i, (n, m12, m21) in zip(kfs, res):
        no, o12, o21 = oracle.search_by_bow(osides[i][0], osides[F][1], 50, 0.75, True)
        assert n == no and np.array_equal(m12, o12) and np.array_equal(m21, o21)
    assert sum(r[0] for r in res) > 100
    # key-frame pairs of different sizes in one batch, with an empty pair in the middle
    empty = orbfe.BowSide(np.zeros((0, 32), np.uint8), np.zeros(0, np.float32), {})
    a = [sides[0][0], empty, sides[3][0], sides[2][0]]
    b = [sides[2][0], sides[1][0], sides[4][0], sides[2][0]]
    res = mt.SearchByBoWBatch(a, b, keyframe_pair=True)
    for (i, j), (n, m12, m21) in zip([(0, 2), (None, 1), (3, 4), (2, 2)], res):
        if i is None:
            assert n == 0 and len(m12) == 0 and (m21 == -1).all()
            continue
        no, o12, o21 = oracle.search_by_bow(osides[i][0], osides[j][0], 49, 0.75, True)
        assert n == no and np.array_equal(m12, o12) and np.array_equal(m21, o21)
    assert mt.SearchByBoWBatch([], []) == []


def test_distinctive_descriptors(orbfe, oracle):
    (k, d, _, _) = _frames(oracle, "mono", 7, 2000)[0]
    rng = np.random.default_rng(8)
    # map points with 0..40 observations (a few with hundreds): noisy copies of one descriptor each, so that medians
    # collide and the first-wins rule matters
    sizes = np.concatenate([rng.integers(0, 41, 3000), [150, 333, 1, 2, 0, 64, 33, 32]]).astype(np.int32)
    start = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    rows = []
    for n in sizes:
        base = d[rng.integers(0, len(d))]
        obs = np.repeat(base[None], n, 0)
        flips = rng.integers(0, 256, (n, 6))
        for j in range(6):
            obs[np.arange(n), flips[:, j] >> 3] ^= (1 << (flips[:, j] & 7)).astype(np.uint8)
        rows.append(obs)
    desc = np.concatenate(rows)
    got = orbfe.ORBmatcher().ComputeDistinctiveDescriptors(desc, start)
    want = oracle.distinctive_descriptors(desc, start)
    assert np.array_equal(got, want)
    assert (got[sizes == 0] == -1).all() and (got[sizes > 0] >= 0).all()
    assert len(orbfe.ORBmatcher().ComputeDistinctiveDescriptors(np.zeros((0, 32), np.uint8), [0])) == 0
    with pytest.raises(orbfe.OrbfeError):
        orbfe.ORBmatcher().ComputeDistinctiveDescriptors(desc, [0, 5, 3])
