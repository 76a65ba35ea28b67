"""GPU parity (through the C ABI) of the key-frame guided searches against the CPU oracle
(oracle/orc_guided.cpp): SearchByProjection(Frame, KeyFrame), SearchByProjection(KeyFrame, Scw), the search of
both Fuse overloads, SearchBySim3 and SearchForTriangulation.  Indices, distances and counts are bit-exact."""
import numpy as np
import pytest

from orb_slam2_annotate_b200 import synth

pytestmark = pytest.mark.gpu

_cache = {}


def _frames(oracle, seed, nf):
    key = (seed, nf)
    if key not in _cache:
        imgs, off = synth.sequence(3, 640, 480, seed=seed)
        out = []
        for im in imgs:
            oc = oracle.OracleExtractor(nf)
            k, d = oc(im)
            out.append((k, d, oc.tables()))
        _cache[key] = (out, off)
    return _cache[key]


def _flip(desc, rng, maxflips):
    desc = desc.copy()
    flips = rng.integers(0, maxflips + 1, len(desc))
    for i in range(len(desc)):
        bits = rng.choice(256, flips[i], replace=False)
        np.bitwise_xor.at(desc[i], bits >> 3, (1 << (bits & 7)).astype(np.uint8))
    return desc


def _points(k, d, n, seed, scale, th, shift=(0.0, 0.0), jitter=4.0, maxflips=60, per_keypoint=False):
    """Map points that project near keypoints of `k`: position = keypoint + shift + jitter, predicted level =
    keypoint octave (+0 / +1), descriptor = the keypoint's with random bit flips."""
    rng = np.random.default_rng(seed)
    src = np.arange(len(k)) if per_keypoint else rng.integers(0, len(k), n)
    n = len(src)
    lvl = np.clip(k["octave"][src] + rng.integers(0, 2, n), 0, 7).astype(np.int32)
    u = (k["x"][src] + shift[0] + rng.uniform(-jitter, jitter, n)).astype(np.float32)
    return dict(valid=(rng.random(n) < 0.9).astype(np.uint8), u=u,
                v=(k["y"][src] + shift[1] + rng.uniform(-jitter, jitter, n)).astype(np.float32),
                ur=(u - 20 + rng.uniform(-2, 2, n)).astype(np.float32),
                radius=(np.float32(th) * scale[lvl]).astype(np.float32), level=lvl,
                angle=(k["angle"][src] + rng.choice(np.array([0.0, 0.0, 0.0, 120.0], np.float32), n)).astype(np.float32),
                desc=_flip(d[src], rng, maxflips))


@pytest.mark.parametrize("th,orb_dist,ori,resident_like", [(10.0, 100, True, False), (6.0, 64, False, False), (15.0, 50, True, False)])
def test_search_by_projection_frame_keyframe(orbfe, oracle, th, orb_dist, ori, resident_like):
    (fr, off) = _frames(oracle, 7, 2000)
    (k, d, t) = fr[1]
    rng = np.random.default_rng(5)
    pts = _points(k, d, 2500, 31, t["scale"], th)
    occ = (rng.random(len(k)) < 0.1).astype(np.uint8)
    F = orbfe.Frame(k, d, 640, 480, t["scale"])
    ng, mg = orbfe.ORBmatcher(0.9, ori).SearchByProjectionKeyFrame(F, pts, orb_dist, occ.copy())
    no, mo = oracle.search_by_projection_keyframe(oracle.frame_view(k, d, 640, 480, t["scale"]), pts, orb_dist, ori, occ)
    assert ng == no and np.array_equal(mg, mo)
    assert ng > 200
    if ori:
        assert (mo == -2).sum() > 0


@pytest.mark.parametrize("th,nmp", [(10.0, 3000), (4.0, 20000)])
def test_search_by_projection_keyframe_sim3(orbfe, oracle, th, nmp):
    (fr, off) = _frames(oracle, 7, 2000)
    (k, d, t) = fr[0]
    rng = np.random.default_rng(6)
    pts = _points(k, d, nmp, 32, t["scale"], th)
    matched = (rng.random(len(k)) < 0.2).astype(np.uint8)
    KF = orbfe.Frame(k, d, 640, 480, t["scale"])
    ng, mg = orbfe.ORBmatcher(0.75, True).SearchByProjectionSim3(KF, pts, matched.copy())
    no, mo = oracle.search_by_projection_sim3(oracle.frame_view(k, d, 640, 480, t["scale"]), pts, matched)
    assert ng == no and np.array_equal(mg, mo)
    assert ng > 200 and (mg[matched == 1] == -1).all()


@pytest.mark.parametrize("gate,stereo,th,nmp", [(False, False, 4.0, 20000), (True, False, 3.0, 20000), (True, True, 3.0, 5000),
                                                (False, True, 20.0, 700)])
def test_fuse_search(orbfe, oracle, gate, stereo, th, nmp):
    (fr, off) = _frames(oracle, 8, 2000)
    (k, d, t) = fr[0]
    rng = np.random.default_rng(7)
    ur = np.where(rng.random(len(k)) < 0.5, k["x"] - 20, -1).astype(np.float32) if stereo else None
    pts = _points(k, d, nmp, 33, t["scale"], th, jitter=2.5)
    sig = t["inv_sigma2"] if gate else None
    KF = orbfe.Frame(k, d, 640, 480, t["scale"], uright=ur)
    bi, bd = orbfe.ORBmatcher().FuseSearch(KF, pts, sig)
    oi, od = oracle.fuse_search(oracle.frame_view(k, d, 640, 480, t["scale"], ur), pts, sig)
    assert np.array_equal(bi, oi) and np.array_equal(bd, od)
    assert (bi >= 0).sum() > 0.2 * nmp and (bi < 0).sum() > 0
    if gate:                                     # the reprojection gate removed candidates the plain search accepts
        pi, pd = oracle.fuse_search(oracle.frame_view(k, d, 640, 480, t["scale"], ur), pts, None)
        assert (pd < od).sum() > 0


@pytest.mark.parametrize("keyframe", [False, True])
def test_keyframe_int_bounds(orbfe, oracle, keyframe):
    """A distorted camera gives fractional image bounds; a KeyFrame truncates them to int for its cell window
    (include/KeyFrame.h:191-194) while its grid is the Frame's: both readings against the oracle."""
    (fr, off) = _frames(oracle, 8, 2000)
    (k, d, t) = fr[0]
    bounds = (-12.7, 655.3, -9.6, 489.4)
    pts = _points(k, d, 20000, 35, t["scale"], 3.0, jitter=2.5)
    KF = orbfe.Frame(k, d, bounds[1], bounds[3], t["scale"], min_x=bounds[0], min_y=bounds[2], keyframe=keyframe)
    fo = oracle.frame_view(k, d, 640, 480, t["scale"], bounds=bounds, keyframe=keyframe)
    bi, bd = orbfe.ORBmatcher().FuseSearch(KF, pts)
    oi, od = oracle.fuse_search(fo, pts)
    assert np.array_equal(bi, oi) and np.array_equal(bd, od)
    matched = np.zeros(len(k), np.uint8)
    ng, mg = orbfe.ORBmatcher().SearchByProjectionSim3(KF, pts, matched.copy())
    no, mo = oracle.search_by_projection_sim3(fo, pts, matched)
    assert ng == no and np.array_equal(mg, mo)


def test_fuse_search_degenerate(orbfe, oracle):
    (fr, off) = _frames(oracle, 8, 2000)
    (k, d, t) = fr[0]
    mt = orbfe.ORBmatcher()
    pts = _points(k, d, 100, 1, t["scale"], 4.0)
    empty = orbfe.Frame(k[:0], d[:0], 640, 480, t["scale"])
    bi, bd = mt.FuseSearch(empty, pts)
    assert (bi == -1).all() and (bd == 256).all()
    none = {kk: v[:0] for kk, v in pts.items()}
    bi, bd = mt.FuseSearch(orbfe.Frame(k, d, 640, 480, t["scale"]), none)
    assert len(bi) == 0
    pts["valid"][:] = 0
    bi, bd = mt.FuseSearch(orbfe.Frame(k, d, 640, 480, t["scale"]), pts)
    assert (bi == -1).all() and (bd == 256).all()
    # windows far outside the image
    pts["valid"][:] = 1; pts["u"][:] = 5000.0
    bi, bd = mt.FuseSearch(orbfe.Frame(k, d, 640, 480, t["scale"]), pts)
    oi, od = oracle.fuse_search(oracle.frame_view(k, d, 640, 480, t["scale"]), pts)
    assert np.array_equal(bi, oi) and np.array_equal(bd, od) and (bi == -1).all()


@pytest.mark.parametrize("th", [7.5, 15.0])
def test_search_by_sim3(orbfe, oracle, th):
    (fr, off) = _frames(oracle, 9, 2000)
    (k1, d1, t), (k2, d2, _) = fr[0], fr[1]
    s12 = (float(off[0][0] - off[1][0]), float(off[0][1] - off[1][1]))      # a KF1 pixel appears at x + s12 in KF2
    p1 = _points(k1, d1, 0, 41, t["scale"], th, shift=s12, jitter=2.0, maxflips=30, per_keypoint=True)
    p2 = _points(k2, d2, 0, 42, t["scale"], th, shift=(-s12[0], -s12[1]), jitter=2.0, maxflips=30, per_keypoint=True)
    KF1 = orbfe.Frame(k1, d1, 640, 480, t["scale"]); KF2 = orbfe.Frame(k2, d2, 640, 480, t["scale"])
    ng, mg = orbfe.ORBmatcher(0.75, True).SearchBySim3(KF1, KF2, p1, p2)
    no, mo = oracle.search_by_sim3(oracle.frame_view(k1, d1, 640, 480, t["scale"]), oracle.frame_view(k2, d2, 640, 480, t["scale"]), p1, p2)
    assert ng == no and np.array_equal(mg, mo)
    assert ng > 100


def _feat_vec(desc, octave, nbits, rng=None, drop=0.0):
    node = (desc[:, 0].astype(np.int64) & ((1 << nbits) - 1)) | (octave.astype(np.int64) << nbits)
    fv = {}
    for i, kk in enumerate(node):
        fv.setdefault(int(kk) * 3 + 1, []).append(i)
    if rng is not None:
        for kk in list(fv):
            if rng.random() < drop:
                del fv[kk]
            else:
                rng.shuffle(fv[kk])
    return fv


@pytest.mark.parametrize("nbits,slope,stereo,only_stereo,ori,epipole", [(2, 0.0, False, False, True, (-2000.0, 240.0)),
                                                                        (0, 0.3, False, False, True, (320.0, 240.0)),
                                                                        (3, -0.7, True, False, False, (100.0, 100.0)),
                                                                        (1, 0.1, True, True, True, (320.0, 240.0))])
def test_search_for_triangulation(orbfe, oracle, nbits, slope, stereo, only_stereo, ori, epipole):
    (fr, off) = _frames(oracle, 9, 2000)
    (k1, d1, t), (k2, d2, _) = fr[0], fr[1]
    dx, dy = float(off[0][0] - off[1][0]), float(off[0][1] - off[1][1])
    # epipolar line of (x1, y1) in image 2: through (x1 + dx, y1 + dy) with the given slope:
    # a = -m, b = 1, c = m*(x1 + dx) - (y1 + dy)
    m = slope
    F12 = np.array([[0, 0, m], [0, 0, -1], [-m, 1, m * dx - dy]], np.float32)
    rng = np.random.default_rng(11)
    has1 = (rng.random(len(k1)) < 0.3).astype(np.uint8); has2 = (rng.random(len(k2)) < 0.3).astype(np.uint8)
    ur1 = np.where(rng.random(len(k1)) < 0.5, k1["x"] - 20, -1).astype(np.float32) if stereo else None
    ur2 = np.where(rng.random(len(k2)) < 0.5, k2["x"] - 20, -1).astype(np.float32) if stereo else None
    fv1 = _feat_vec(d1, k1["octave"], nbits, rng, 0.1); fv2 = _feat_vec(d2, k2["octave"], nbits, rng, 0.1)
    S1 = orbfe.TriSide(k1, d1, has1, fv1, ur1); S2 = orbfe.TriSide(k2, d2, has2, fv2, ur2)
    ng, mg = orbfe.ORBmatcher(0.6, ori).SearchForTriangulation(S1, S2, F12, epipole[0], epipole[1], t["scale"], t["sigma2"], only_stereo)
    no, mo = oracle.search_for_triangulation(oracle.tri_side(k1, d1, has1, fv1, ur1), oracle.tri_side(k2, d2, has2, fv2, ur2), F12,
                                             epipole[0], epipole[1], t["scale"], t["sigma2"], only_stereo, ori)
    assert ng == no and np.array_equal(mg, mo)
    assert ng > (20 if only_stereo else 60)
    assert (mg[has1 == 1] == -1).all()


def test_search_for_triangulation_degenerate(orbfe, oracle):
    (fr, off) = _frames(oracle, 9, 2000)
    (k1, d1, t), (k2, d2, _) = fr[0], fr[1]
    mt = orbfe.ORBmatcher(0.6, True)
    F12 = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32)
    z1 = np.zeros(len(k1), np.uint8); z2 = np.zeros(len(k2), np.uint8)
    fv1 = _feat_vec(d1, k1["octave"], 2); fv2 = _feat_vec(d2, k2["octave"], 2)
    S1 = orbfe.TriSide(k1, d1, z1, fv1)
    # empty second key frame, empty feature vector, disjoint node ids
    n, mm = mt.SearchForTriangulation(S1, orbfe.TriSide(k2[:0], d2[:0], z2[:0], {}), F12, 0, 0, t["scale"], t["sigma2"])
    assert n == 0 and (mm == -1).all()
    n, mm = mt.SearchForTriangulation(S1, orbfe.TriSide(k2, d2, z2, {}), F12, 0, 0, t["scale"], t["sigma2"])
    assert n == 0 and (mm == -1).all()
    n, mm = mt.SearchForTriangulation(S1, orbfe.TriSide(k2, d2, z2, {kk + 1: v for kk, v in fv2.items()}), F12, 0, 0, t["scale"], t["sigma2"])
    assert n == 0 and (mm == -1).all()
    # a degenerate fundamental matrix (den == 0 for every point, :167) matches nothing
    n, mm = mt.SearchForTriangulation(S1, orbfe.TriSide(k2, d2, z2, fv2), np.zeros((3, 3), np.float32), 0, 0, t["scale"], t["sigma2"])
    no, mo = oracle.search_for_triangulation(oracle.tri_side(k1, d1, z1, fv1), oracle.tri_side(k2, d2, z2, fv2), np.zeros(9, np.float32),
                                             0, 0, t["scale"], t["sigma2"], False, True)
    assert n == no == 0 and np.array_equal(mm, mo)
    # the same key frame against itself with the identity-shift geometry: every free feature finds itself or a twin
    n, mm = mt.SearchForTriangulation(S1, orbfe.TriSide(k1, d1, z1, fv1), F12, -5000.0, 0.0, t["scale"], t["sigma2"])
    no, mo = oracle.search_for_triangulation(oracle.tri_side(k1, d1, z1, fv1), oracle.tri_side(k1, d1, z1, fv1), F12, -5000.0, 0.0,
                                             t["scale"], t["sigma2"], False, True)
    assert n == no and np.array_equal(mm, mo) and n > 0.8 * len(k1)
