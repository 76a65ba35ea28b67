"""Known-answer tests for the oracle's matcher restatements (oracle/orc_matcher.cpp), derived by
hand from the cited reference lines.  (The same routines are compared with the unmodified reference TUs in
tests/test_oracle_vs_ref_matcher.py.)"""
import numpy as np

from oracle.oracle_py import KP_DTYPE

SCALE = (1.2 ** np.arange(8)).astype(np.float32)


def kp(x, y, octave=0, angle=0.0):
    return (x, y, 31.0, angle, 50.0, octave, -1)


def test_descriptor_distance_kats(oracle):
    z = np.zeros(32, np.uint8); o = np.full(32, 255, np.uint8)
    assert oracle.descriptor_distance(z, z) == 0 and oracle.descriptor_distance(z, o) == 256      # ORBmatcher.cc:1828-1844
    for bit in (0, 7, 8, 100, 255):
        a = z.copy(); a[bit >> 3] = 1 << (bit & 7)
        assert oracle.descriptor_distance(a, z) == 1
    rng = np.random.default_rng(0)
    for _ in range(100):
        a = rng.integers(0, 256, 32, dtype=np.uint8); b = rng.integers(0, 256, 32, dtype=np.uint8)
        assert oracle.descriptor_distance(a, b) == int(np.unpackbits(a ^ b).sum())


def test_three_maxima(oracle):
    import ctypes as C
    def run(counts):
        c = np.zeros(30, np.int32); c[:len(counts)] = counts
        a, b, d = C.c_int(), C.c_int(), C.c_int()
        oracle.lib().orc_three_maxima(c.ctypes.data, 30, C.byref(a), C.byref(b), C.byref(d))
        return a.value, b.value, d.value
    assert run([5, 9, 7]) == (1, 2, 0)
    assert run([100, 9, 7]) == (0, -1, -1)          # second < 10% of first drops second and third (:1808-1812)
    assert run([100, 50, 7]) == (0, 1, -1)          # third < 10% of first (:1813-1816)
    assert run([3, 3, 3]) == (0, 1, 2)              # strict '>' keeps the first of equals in front


def test_grid_round_vs_floor_quirk(oracle):
    # PosInGrid rounds (Frame.cc:422-423): x=14.9 at 10 px cells lands in column 1, and a query
    # window [floor, ceil] still finds it
    k = np.array([kp(14.9, 5.0), kp(300.0, 200.0, 1)], KP_DTYPE)
    d = np.zeros((2, 32), np.uint8)
    fv = oracle.frame_view(k, d, 640, 480, SCALE)
    cs = np.zeros(3073, np.int32); ci = np.zeros(2, np.int32)
    oracle.lib().orc_assign_grid(fv[0], cs.ctypes.data, ci.ctypes.data)
    cell = 1 * 48 + 1    # round(14.9/10)=1, round(5.0/10)=1 (round-half-away: 0.5 -> 1)
    assert cs[cell + 1] - cs[cell] == 1
    out = np.zeros(4, np.int32)
    import ctypes as C
    n = oracle.lib().orc_features_in_area(C.byref(fv[0]), cs.ctypes.data, ci.ctypes.data, 12.0, 5.0, 5.0, -1, -1, out.ctypes.data, 4)
    assert n == 1 and out[0] == 0
    # level filter: minLevel=0,maxLevel=0 checks levels (maxLevel>=0), octave-1 keypoint excluded (Frame.cc:381-395)
    n = oracle.lib().orc_features_in_area(C.byref(fv[0]), cs.ctypes.data, ci.ctypes.data, 300.0, 200.0, 5.0, 0, 0, out.ctypes.data, 4)
    assert n == 0
    # strict '<' on the window: a keypoint exactly r away is excluded (Frame.cc:405)
    n = oracle.lib().orc_features_in_area(C.byref(fv[0]), cs.ctypes.data, ci.ctypes.data, 295.0, 200.0, 5.0, -1, -1, out.ctypes.data, 4)
    assert n == 0


def test_search_for_initialization_kat(oracle):
    # F1: two octave-0 keypoints with the same descriptor both near F2[0]; the second query's
    # candidate is skipped because vMatchedDistance[i2] <= dist (:515-516) -> only the first matches.
    d0 = np.zeros(32, np.uint8)
    far = np.full(32, 255, np.uint8)
    k1 = np.array([kp(100, 100), kp(102, 100), kp(400, 300, 1)], KP_DTYPE)
    d1 = np.stack([d0, d0, d0])
    k2 = np.array([kp(101, 100), kp(110, 100)], KP_DTYPE)
    d2 = np.stack([d0, far])
    prev = np.ascontiguousarray(np.stack([k1["x"], k1["y"]], 1), np.float32)
    n, m12 = oracle.search_for_initialization(oracle.frame_view(k1, d1, 640, 480, SCALE),
                                              oracle.frame_view(k2, d2, 640, 480, SCALE), prev, 100, 0.9, False)
    assert n == 1 and m12.tolist() == [0, -1, -1]
    assert prev[0].tolist() == [101.0, 100.0] and prev[1].tolist() == [102.0, 100.0]     # :595-597
    # a strictly better later query steals the match and un-matches the earlier one (:536-541)
    d1b = d1.copy(); d1b[0, 0] = 0b111           # query 0 at distance 3, query 1 at distance 0
    prev = np.ascontiguousarray(np.stack([k1["x"], k1["y"]], 1), np.float32)
    n, m12 = oracle.search_for_initialization(oracle.frame_view(k1, d1b, 640, 480, SCALE),
                                              oracle.frame_view(k2, d2, 640, 480, SCALE), prev, 100, 0.9, False)
    assert n == 1 and m12.tolist() == [-1, 0, -1]
    # ratio test: best must be < ratio * second (:534)
    d2c = np.stack([d0, d0]); d2c[1, 0] = 1      # distances 0 and 1 -> 0 < 0.9*1 passes
    n, m12 = oracle.search_for_initialization(oracle.frame_view(k1[:1], d1[:1], 640, 480, SCALE),
                                              oracle.frame_view(k2, d2c, 640, 480, SCALE),
                                              np.array([[100, 100]], np.float32), 100, 0.9, False)
    assert n == 1 and m12.tolist() == [0]
    d1d = d1[:1].copy(); d1d[0, 1] = 0xFF         # distances 8 and 9 -> 8 < 8.1 passes; with ratio 0.8 fails
    n, _ = oracle.search_for_initialization(oracle.frame_view(k1[:1], d1d, 640, 480, SCALE),
                                            oracle.frame_view(k2, d2c, 640, 480, SCALE),
                                            np.array([[100, 100]], np.float32), 100, 0.8, False)
    assert n == 0


def test_search_by_projection_kat(oracle):
    d0 = np.zeros(32, np.uint8)
    k = np.array([kp(100, 100, 0), kp(103, 100, 0)], KP_DTYPE)
    d = np.stack([d0, d0]); d[1, 0] = 0xFF     # distance 8 from the map points
    f = oracle.frame_view(k, d, 640, 480, SCALE)
    mps = dict(track_in_view=[1, 1, 0], proj_x=[100, 100, 100], proj_y=[100, 100, 100], proj_xr=[0, 0, 0], level=[0, 0, 0],
               view_cos=[0.9, 0.9, 0.9], desc=np.stack([d0, d0, d0]), observations=[1, 1, 1])
    occ = np.zeros(2, np.uint8)
    n, out = oracle.search_by_projection(f, mps, 1.0, 0.8, occ)
    # r = 4.0 (viewCos <= 0.998), both keypoints inside; same octave and 0 > 0.8*8 false -> accept kp0 (:127-135);
    # second map point finds kp0 occupied (:95-98), best becomes kp1 at distance 8 with no second -> accepted
    assert n == 2 and out.tolist() == [0, 1] and occ.tolist() == [1, 1]
    # viewCos > 0.998 -> r = 2.5: kp1 (3 px away) is outside
    mps["view_cos"] = [0.999, 0.999, 0.999]
    occ = np.zeros(2, np.uint8)
    n, out = oracle.search_by_projection(f, mps, 1.0, 0.8, occ)
    assert n == 1 and out.tolist() == [0, -1]


def test_search_by_projection_frame_kat(oracle):
    """SearchByProjection(CurrentFrame, LastFrame) known answers (ORBmatcher.cc:1484-1630)."""
    d0 = np.zeros(32, np.uint8)
    # two keypoints 3 px apart, both octave 1; angles 0 and 100 degrees
    k = np.array([kp(100, 100, 1, 0.0), kp(103, 100, 1, 100.0)], KP_DTYPE)
    d = np.stack([d0, d0]); d[1, 0] = 0xFF                      # kp1 is 8 bits away from the points
    f = oracle.frame_view(k, d, 640, 480, SCALE)

    def pts(n, obs, minl, maxl, angle, valid=None, radius=5.0):
        return dict(valid=valid or [1] * n, u=[100] * n, v=[100] * n, ur=[0] * n, radius=[radius] * n,
                    min_level=[minl] * n, max_level=[maxl] * n, angle=angle, desc=np.stack([d0] * n), observations=obs)

    occ = np.zeros(2, np.uint8)
    # point 0 has no observations: it takes kp0 (distance 0) but does not block it, so point 1 takes kp0 again
    # (two match events, :1587-1588), point 2 (observations) takes it a third time; no orientation check
    n, out = oracle.search_by_projection_frame(f, pts(3, [0, 0, 2], 0, 2, [0, 0, 0]), False, occ)
    assert n == 3 and out.tolist() == [2, -1]
    # with observations the first point blocks kp0 (:1567-1569): the second falls back to kp1
    n, out = oracle.search_by_projection_frame(f, pts(2, [1, 1], 0, 2, [0, 0]), False, occ)
    assert n == 2 and out.tolist() == [0, 1]
    # level window (:1547-1552): octave-1 keypoints are outside (2, -1) = "octave >= 2"
    n, out = oracle.search_by_projection_frame(f, pts(1, [1], 2, -1, [0]), False, occ)
    assert n == 0 and out.tolist() == [-1, -1]
    # invalid points and initially occupied keypoints are skipped
    n, out = oracle.search_by_projection_frame(f, pts(2, [1, 1], 0, 2, [0, 0], valid=[0, 1]), False, np.array([1, 0], np.uint8))
    assert n == 1 and out.tolist() == [-1, 1]
    # rotation check (:1611-1628): three events in bin 0 (rot 0) and one in bin round(260/30) = 9; bin 9 holds
    # 1 < 0.1 * 3 ... no: 1 >= 0.3, so it survives as second maximum; make it 11 vs 1 to drop it
    k2 = np.array([kp(100 + 8 * i, 100, 1, 0.0) for i in range(11)] + [kp(300, 100, 1, 100.0)], KP_DTYPE)
    f2 = oracle.frame_view(k2, np.stack([d0] * 12), 640, 480, SCALE)
    p = dict(valid=[1] * 12, u=[100 + 8 * i for i in range(11)] + [300], v=[100] * 12, ur=[0] * 12, radius=[2.0] * 12,
             min_level=[0] * 12, max_level=[2] * 12, angle=[0.0] * 12, desc=np.stack([d0] * 12), observations=[1] * 12)
    n, out = oracle.search_by_projection_frame(f2, p, True, np.zeros(12, np.uint8))
    assert n == 11 and out.tolist() == list(range(11)) + [-2]
    n, out = oracle.search_by_projection_frame(f2, p, False, np.zeros(12, np.uint8))
    assert n == 12 and out.tolist() == list(range(12))


def test_knn2_kat(oracle):
    t = np.zeros((3, 32), np.uint8); t[1, 0] = 1; t[2, 0] = 3
    q = np.zeros((1, 32), np.uint8)
    idx, d1, d2 = oracle.knn2(q, t, 50, 0.7)
    assert (idx[0], d1[0], d2[0]) == (0, 0, 1)
    t[0, 0] = 1        # two rows at distance 1: first wins, d2 == d1, ratio test rejects
    idx, d1, d2 = oracle.knn2(q, t, 50, 0.7)
    assert (idx[0], d1[0], d2[0]) == (-1, 1, 1)


def _bits(k):
    """descriptor with the first k bits set: Hamming distance k from the zero descriptor"""
    d = np.zeros(32, np.uint8)
    for b in range(k):
        d[b >> 3] |= 1 << (b & 7)
    return d


def test_search_by_bow_kat(oracle):
    """Hand-derived cases for SearchByBoW (ORBmatcher.cc:185-325 and :610-743)."""
    z = _bits(0)
    # side 1: three features, all zero descriptors; nodes {1: [0], 5: [1, 2]}.  side 2: nodes {5: [0, 1, 2], 7: [3]}
    d2 = np.stack([_bits(3), _bits(10), _bits(30), _bits(0)])
    s1 = oracle.bow_view(np.stack([z, z, z]), [0, 0, 0], {1: [0], 5: [1, 2]})
    s2 = oracle.bow_view(d2, [0, 0, 0, 0], {5: [0, 1, 2], 7: [3]})
    # node 1 has no partner (lower_bound skips it, :289); in node 5 feature 1 takes side-2 feature 0
    # (3 < 0.7 * 10), then feature 2 sees only {1, 2}: 10 < 0.7 * 30 -> takes 1.  Side-2 feature 3 (node 7,
    # distance 0) is never looked at.
    n, m12, m21 = oracle.search_by_bow(s1, s2, 50, 0.7, False)
    assert n == 2 and m12.tolist() == [-1, 0, 1] and m21.tolist() == [1, 2, -1, -1]
    # ratio 0.25: 3 < 2.5 fails for feature 1, so side-2 feature 0 stays free and feature 2 fails the same way
    n, m12, m21 = oracle.search_by_bow(s1, s2, 50, 0.25, False)
    assert n == 0 and (m12 == -1).all() and (m21 == -1).all()
    # threshold: best == 50 passes with th = 50 (<= TH_LOW, :263) and fails with th = 49 (< TH_LOW, :686)
    a = oracle.bow_view(np.stack([z]), [0], {3: [0]})
    b = oracle.bow_view(np.stack([_bits(50), _bits(200)]), [0, 0], {3: [0, 1]})
    assert oracle.search_by_bow(a, b, 50, 0.7, False)[0] == 1
    assert oracle.search_by_bow(a, b, 49, 0.7, False)[0] == 0
    # two candidates at the same distance: the first is best, the second becomes second-best (:253-256),
    # and d < ratio * d fails
    b = oracle.bow_view(np.stack([_bits(8), _bits(8)]), [0, 0], {3: [0, 1]})
    assert oracle.search_by_bow(a, b, 50, 0.9, False)[0] == 0
    # a lone candidate: second-best stays 256
    b = oracle.bow_view(np.stack([_bits(40)]), [0], {3: [0]})
    n, m12, m21 = oracle.search_by_bow(a, b, 50, 0.7, False)
    assert n == 1 and m12.tolist() == [0] and m21.tolist() == [0]
    # map point flags: an invalid side-1 feature is skipped (:214-220), an invalid side-2 feature is not a
    # candidate in the KeyFrame-pair form (:662-670)
    s1 = oracle.bow_view(np.stack([z, z]), [0, 0], {2: [0, 1]}, valid=[0, 1])
    s2 = oracle.bow_view(np.stack([_bits(1), _bits(20)]), [0, 0], {2: [0, 1]}, valid=[0, 1])
    n, m12, m21 = oracle.search_by_bow(s1, s2, 49, 0.7, False)
    assert n == 1 and m12.tolist() == [-1, 1] and m21.tolist() == [-1, 1]
    # feature order inside a node is the vector's order, not the index order
    s1 = oracle.bow_view(np.stack([z, z]), [0, 0], {2: [1, 0]})
    s2 = oracle.bow_view(np.stack([_bits(2), _bits(12)]), [0, 0], {2: [1, 0]})
    n, m12, m21 = oracle.search_by_bow(s1, s2, 50, 0.7, False)
    assert n == 2 and m12.tolist() == [1, 0]        # feature 1 goes first and takes side-2 feature 0
    # rotation check (:298-318): eleven matches with rot 0 (bin 0) and one with rot 100 (bin round(100/30) = 3):
    # 1 < 0.1 * 11, so bin 3 is not a maximum and its match is removed from both result arrays
    n1 = 12
    d1 = np.stack([_bits(0)] * n1)
    d2 = np.stack([_bits(1)] * n1)
    fv = {i: [i] for i in range(n1)}
    s1 = oracle.bow_view(d1, [0.0] * 11 + [100.0], fv)
    s2 = oracle.bow_view(d2, [0.0] * n1, fv)
    n, m12, m21 = oracle.search_by_bow(s1, s2, 50, 0.7, True)
    assert n == 11 and m12.tolist() == list(range(11)) + [-1] and m21.tolist() == list(range(11)) + [-1]
    n, m12, m21 = oracle.search_by_bow(s1, s2, 50, 0.7, False)
    assert n == 12 and m12.tolist() == list(range(12))


def test_distinctive_descriptors_kat(oracle):
    """MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:266-334)."""
    # three descriptors at 0, 10 and 30 bits: rows sorted = [0,10,30], [0,10,20], [0,20,30]; medians (index 1)
    # 10, 10, 20 -> the first of the two least medians wins (strict '<', :317)
    d = np.stack([_bits(0), _bits(10), _bits(30)])
    assert oracle.distinctive_descriptors(d, [0, 3]).tolist() == [0]
    # even N takes the lower middle: index (int)(0.5*(4-1)) = 1.  Rows of {0,4,6,40}: [0,4,6,40] -> 4,
    # [0,2,4,36] -> 2, [0,2,6,34] -> 2, [0,34,36,40] -> 34: descriptor 1 (first with median 2)
    d4 = np.stack([_bits(0), _bits(4), _bits(6), _bits(40)])
    assert oracle.distinctive_descriptors(d4, [0, 4]).tolist() == [1]
    # a single observation, an empty set, and several points in one batch
    both = np.concatenate([d, d4, d[:1]])
    assert oracle.distinctive_descriptors(both, [0, 3, 7, 7, 8]).tolist() == [0, 1, -1, 0]
    # two observations: medians are the zero diagonals -> the first
    assert oracle.distinctive_descriptors(d[1:], [0, 2]).tolist() == [0]
