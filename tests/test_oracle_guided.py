"""Known-answer tests for oracle/orc_guided.cpp (the remaining guided searches of ORBmatcher), derived by
hand from the cited reference lines.  (The same routines are compared with the unmodified reference TU in
tests/test_oracle_vs_ref_matcher.py.)"""
import numpy as np

from oracle.oracle_py import KP_DTYPE

SCALE = (1.2 ** np.arange(8)).astype(np.float32)
SIGMA2 = (SCALE * SCALE).astype(np.float32)
INV_SIGMA2 = (np.float32(1.0) / SIGMA2).astype(np.float32)


def kp(x, y, octave=0, angle=0.0):
    return (x, y, 31.0, angle, 50.0, octave, -1)


def bits(k):
    """descriptor with the k lowest bits set"""
    d = np.zeros(32, np.uint8)
    for b in range(k):
        d[b >> 3] |= 1 << (b & 7)
    return d


def pts(n, level, u=100.0, v=100.0, radius=5.0, valid=None, angle=None, ur=None, desc=None):
    return dict(valid=valid or [1] * n, u=[u] * n, v=[v] * n, ur=ur or [0.0] * n, radius=[radius] * n, level=level,
                angle=angle or [0.0] * n, desc=desc if desc is not None else np.stack([bits(0)] * n))


def test_sbp_frame_keyframe_kat(oracle):
    """SearchByProjection(Frame&, KeyFrame*, ...) — src/ORBmatcher.cc:1641-1770."""
    k = np.array([kp(100, 100, 1), kp(103, 100, 2), kp(101, 100, 3)], KP_DTYPE)
    d = np.stack([bits(0), bits(8), bits(0)])
    f = oracle.frame_view(k, d, 640, 480, SCALE)
    occ = np.zeros(3, np.uint8)
    # any assigned keypoint is taken (:1720): two points of level 1 get kp0 (dist 0) then kp1 (dist 8, octave 2 <= 1+1);
    # kp2 (octave 3) is outside the level window [0, 2]
    n, out = oracle.search_by_projection_keyframe(f, pts(3, [1, 1, 1]), 100, False, occ)
    assert n == 2 and out.tolist() == [0, 1, -1]
    # ORBdist bounds the accepted distance (:1734): with ORBdist 7 the second point finds nothing
    n, out = oracle.search_by_projection_keyframe(f, pts(2, [1, 1]), 7, False, occ)
    assert n == 1 and out.tolist() == [0, -1, -1]
    # level window is symmetric here (nPredictedLevel-1 .. +1, :1705): level 3 sees kp1 (octave 2) and kp2 (octave 3)
    n, out = oracle.search_by_projection_keyframe(f, pts(1, [3]), 100, False, occ)
    assert n == 1 and out.tolist() == [-1, -1, 0]
    # keypoints that already hold a map point are skipped whatever its observations
    n, out = oracle.search_by_projection_keyframe(f, pts(1, [1]), 100, False, np.array([1, 0, 0], np.uint8))
    assert n == 1 and out.tolist() == [-1, 0, -1]
    # rotation check: 11 matches at rot 0 and one at rot 260 -> bin 9 holds 1 < 0.1*11: cleared (-2)
    k2 = np.array([kp(100 + 8 * i, 100, 1, 0.0) for i in range(11)] + [kp(300, 100, 1, 100.0)], KP_DTYPE)
    f2 = oracle.frame_view(k2, np.stack([bits(0)] * 12), 640, 480, SCALE)
    p = dict(valid=[1] * 12, u=[100 + 8 * i for i in range(11)] + [300], v=[100] * 12, radius=[2.0] * 12, level=[1] * 12,
             angle=[0.0] * 12, desc=np.stack([bits(0)] * 12))
    n, out = oracle.search_by_projection_keyframe(f2, p, 100, True, np.zeros(12, np.uint8))
    assert n == 11 and out.tolist() == list(range(11)) + [-2]


def test_sbp_keyframe_sim3_kat(oracle):
    """SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th) — src/ORBmatcher.cc:335-451."""
    k = np.array([kp(100, 100, 1), kp(103, 100, 2), kp(101, 100, 0)], KP_DTYPE)
    d = np.stack([bits(0), bits(8), bits(60)])
    f = oracle.frame_view(k, d, 640, 480, SCALE)
    none = np.zeros(3, np.uint8)
    # level window is [nPredictedLevel-1, nPredictedLevel] (:426): level 1 sees kp0 (octave 1) and kp2 (octave 0),
    # kp2 is 60 bits away > TH_LOW (:440): first point takes kp0, second finds only kp2 -> rejected
    n, out = oracle.search_by_projection_sim3(f, pts(2, [1, 1]), none)
    assert n == 1 and out.tolist() == [0, -1, -1]
    # level 2 sees kp0 and kp1; vpMatched[0] set on entry (:421): the point takes kp1 (8 bits)
    n, out = oracle.search_by_projection_sim3(f, pts(1, [2]), np.array([1, 0, 0], np.uint8))
    assert n == 1 and out.tolist() == [-1, 0, -1]
    # distance exactly TH_LOW is accepted (<=)
    d2 = d.copy(); d2[2] = bits(50)
    f2 = oracle.frame_view(k, d2, 640, 480, SCALE)
    n, out = oracle.search_by_projection_sim3(f2, pts(1, [0]), none)
    assert n == 1 and out.tolist() == [-1, -1, 0]
    d2[2] = bits(51)
    n, out = oracle.search_by_projection_sim3(oracle.frame_view(k, d2, 640, 480, SCALE), pts(1, [0]), none)
    assert n == 0


def test_fuse_search_kat(oracle):
    """Fuse(pKF, vpMapPoints, th) :940-1090 and Fuse(pKF, Scw, ...) :1100-1237 — the search."""
    k = np.array([kp(100, 100, 1), kp(104, 100, 1), kp(100, 101, 1)], KP_DTYPE)
    d = np.stack([bits(5), bits(0), bits(3)])
    f = oracle.frame_view(k, d, 640, 480, SCALE)
    # Sim3 variant: no reprojection gate, every point is independent -> both points choose kp1 (distance 0)
    bi, bd = oracle.fuse_search(f, pts(2, [1, 1], radius=6.0))
    assert bi.tolist() == [1, 1] and bd.tolist() == [0, 0]
    # pose variant, monocular gate (:1033-1041): e2 * invSigma2[1] > 5.99.  kp1 is 4 px away: 16/1.44 = 11.1 > 5.99
    # -> dropped; kp2 is 1 px away -> stays (3 bits), kp0 at 0 px stays (5 bits): best = kp2
    bi, bd = oracle.fuse_search(f, pts(1, [1], radius=6.0), INV_SIGMA2)
    assert bi.tolist() == [2] and bd.tolist() == [3]
    # stereo gate (:1020-1031): threshold 7.8 with the right coordinate's error.  kp1: ex 4 -> 16/1.44 > 7.8 dropped;
    # kp2 with er = 3: (1 + 9)/1.44 = 6.94 <= 7.8 stays; a keypoint with uright < 0 keeps the 5.99 gate
    ur = np.array([-1.0, 50.0, 50.0], np.float32)
    fs = oracle.frame_view(k, d, 640, 480, SCALE, uright=ur)
    bi, bd = oracle.fuse_search(fs, pts(1, [1], radius=6.0, ur=[53.0]), INV_SIGMA2)
    assert bi.tolist() == [2] and bd.tolist() == [3]
    bi, bd = oracle.fuse_search(fs, pts(1, [1], radius=6.0, ur=[53.5]), INV_SIGMA2)     # (1 + 12.25)/1.44 = 9.2 > 7.8
    assert bi.tolist() == [0] and bd.tolist() == [5]
    # best distance above TH_LOW: no fusion, the distance is still reported
    d3 = np.stack([bits(51)] * 3)
    bi, bd = oracle.fuse_search(oracle.frame_view(k, d3, 640, 480, SCALE), pts(1, [1], radius=6.0))
    assert bi.tolist() == [-1] and bd.tolist() == [51]
    # level window [l-1, l] (:1016): level 3 sees nothing of octave 1; invalid points give -1 / 256
    bi, bd = oracle.fuse_search(f, pts(2, [3, 1], radius=6.0, valid=[1, 0]))
    assert bi.tolist() == [-1, -1] and bd.tolist() == [256, 256]


def test_search_by_sim3_kat(oracle):
    """SearchBySim3 — src/ORBmatcher.cc:1251-1477: two independent searches, kept where they agree."""
    k1 = np.array([kp(100, 100, 1), kp(200, 100, 1)], KP_DTYPE)
    k2 = np.array([kp(300, 300, 1), kp(400, 300, 1), kp(302, 300, 1)], KP_DTYPE)
    d1 = np.stack([bits(0), bits(20)])
    d2 = np.stack([bits(0), bits(20), bits(2)])
    f1 = oracle.frame_view(k1, d1, 640, 480, SCALE); f2 = oracle.frame_view(k2, d2, 640, 480, SCALE)
    # KF1 points (descriptors = KF1's) projected into KF2: point 0 -> (300,300), point 1 -> (400,300)
    p1 = dict(valid=[1, 1], u=[300, 400], v=[300, 300], radius=[5.0, 5.0], level=[1, 1], desc=d1)
    # KF2 points into KF1: point 0 -> (100,100), point 1 -> (200,100), point 2 -> (100,100) as well
    p2 = dict(valid=[1, 1, 1], u=[100, 200, 100], v=[100, 100, 100], radius=[5.0] * 3, level=[1] * 3, desc=d2)
    n, m = oracle.search_by_sim3(f1, f2, p1, p2)
    assert n == 2 and m.tolist() == [0, 1]
    # disagreement: KF1 point 0 now prefers kp2 of KF2 (distance 0 vs 2), whose own search returns kp0 of KF1 — equal
    # index, so it still agrees; but make KF2's point 2 invalid and the pair is dropped (:1452)
    d2b = d2.copy(); d2b[0] = bits(9)
    f2b = oracle.frame_view(k2, d2b, 640, 480, SCALE)
    p2b = dict(p2, desc=d2b, valid=[1, 1, 0])
    n, m = oracle.search_by_sim3(f1, f2b, dict(p1), p2b)
    assert n == 1 and m.tolist() == [-1, 1]
    # TH_HIGH bound (:1361): a best distance of 101 bits gives no match
    p1c = dict(p1, desc=np.stack([bits(103), bits(20)]))   # 103 / 101 bits from KF2's kp0 / kp2
    n, m = oracle.search_by_sim3(f1, f2, p1c, p2)
    assert n == 1 and m.tolist() == [-1, 1]


def test_search_for_triangulation_kat(oracle):
    """SearchForTriangulation — src/ORBmatcher.cc:754-928 with CheckDistEpipolarLine :156-175."""
    # F12 such that the epipolar line of kp1 = (x, y) in image 2 is the horizontal line y2 = y:
    # a = 0, b = 1, c = -y  <=>  F12 = [[0,0,0],[0,0,-1],[0,1,0]]  (row-major; a = x*F00 + y*F10 + F20 ...)
    F12 = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32)
    k1 = np.array([kp(100, 100, 0, 10.0), kp(150, 200, 0, 10.0)], KP_DTYPE)
    k2 = np.array([kp(300, 100, 0, 10.0), kp(310, 101.5, 0, 10.0), kp(320, 100.5, 1, 10.0), kp(330, 200, 0, 10.0)], KP_DTYPE)
    d1 = np.stack([bits(0), bits(0)])
    d2 = np.stack([bits(4), bits(2), bits(4), bits(50)])
    none1 = np.zeros(2, np.uint8); none2 = np.zeros(4, np.uint8)
    fv1 = {7: [0, 1]}; fv2 = {7: [0, 1, 2, 3]}
    s1 = oracle.tri_side(k1, d1, none1, fv1)
    far = (-1000.0, -1000.0)     # epipole far away: the parallax gate (:845-853) never fires

    def run(s2, ex=far[0], ey=far[1], **kw):
        return oracle.search_for_triangulation(s1, s2, F12, ex, ey, SCALE, SIGMA2, **kw)

    # feature 0 (y = 100): kp0 of KF2 on the line (dist 4), kp1 1.5 px off: 2.25 < 3.84*1 -> passes with 2 bits and
    # becomes best; kp2 is 0.5 px off with 4 bits > bestDist 2 -> skipped (:840).  feature 1 (y = 200): kp3 with
    # exactly TH_LOW = 50 bits passes 'dist>TH_LOW' (:840).
    n, m = run(oracle.tri_side(k2, d2, none2, fv2), check_ori=False)
    assert n == 2 and m.tolist() == [1, 3]
    # equal distances: the later candidate replaces the earlier one ('dist>bestDist' is strict, :840)
    d2e = d2.copy(); d2e[1] = bits(4)
    n, m = run(oracle.tri_side(k2, d2e, none2, fv2), check_ori=False)
    assert m.tolist()[0] == 2
    # 2 px off the line: at octave 0 4 > 3.84 fails (kp1, 2 bits), at octave 1 sigma2 = 1.44: 4 < 5.53 passes (:172),
    # and kp2's 4 bits tie with kp0's, which ':840' lets through: kp2 replaces kp0
    k2o = k2.copy(); k2o["y"][1] = 102.0; k2o["y"][2] = 102.0
    n, m = run(oracle.tri_side(k2o, d2, none2, fv2), check_ori=False)
    assert m.tolist()[0] == 2
    k2o["octave"][2] = 0                           # the same offset at octave 0 fails: kp0 stays
    n, m = run(oracle.tri_side(k2o, d2, none2, fv2), check_ori=False)
    assert m.tolist()[0] == 0
    # key points of KF2 holding a map point are skipped (:826-828); KF1 features with a map point too (:799-803)
    n, m = run(oracle.tri_side(k2, d2, np.array([0, 1, 0, 0], np.uint8), fv2), check_ori=False)
    assert m.tolist() == [2, 3]                    # kp0 (4 bits) then kp2 (4 bits, 0.5 px): later equal wins
    s1b = oracle.tri_side(k1, d1, np.array([1, 0], np.uint8), fv1)
    n, m = oracle.search_for_triangulation(s1b, oracle.tri_side(k2, d2, none2, fv2), F12, far[0], far[1], SCALE, SIGMA2,
                                           check_ori=False)
    assert n == 1 and m.tolist() == [-1, 3]
    # parallax gate (:845-853): epipole at (300, 100) is 0 px from kp0, 10.1 px from kp1 (102.25 >= 100 passes),
    # 20 px from kp2: kp0 is dropped, kp1 stays best
    n, m = run(oracle.tri_side(k2, d2e, none2, fv2), ex=300.0, ey=100.0, check_ori=False)
    assert m.tolist()[0] == 2
    k2p = k2.copy(); k2p["x"][1] = 309.0            # 9.1 px from the epipole: 83 < 100 -> dropped
    k2p["x"][2] = 305.0                             # octave 1: 25.25 < 120 -> dropped too
    n, m = run(oracle.tri_side(k2p, d2e, none2, fv2), ex=300.0, ey=100.0, check_ori=False)
    assert m.tolist() == [-1, 3]
    # only-stereo (:807-809, :832-834): monocular sides never match
    n, m = run(oracle.tri_side(k2, d2, none2, fv2), only_stereo=True, check_ori=False)
    assert n == 0
    # different vocabulary nodes never meet
    n, m = run(oracle.tri_side(k2, d2, none2, {8: [0, 1, 2, 3]}), check_ori=False)
    assert n == 0 and m.tolist() == [-1, -1]
    # rotation check: both matches in bin 0 survive
    n, m = run(oracle.tri_side(k2, d2, none2, fv2), check_ori=True)
    assert n == 2 and m.tolist() == [1, 3]


def test_keyframe_int_bounds_kat(oracle):
    """KeyFrame::GetFeaturesInArea subtracts int-truncated bounds (KeyFrame.h:191-194, KeyFrame.cc:616-628) while the
    grid is the Frame's.  The truncation moves the cell window by less than a pixel and the window has half a cell of
    slack around every keypoint within the radius (PosInGrid rounds, the window floors / ceils), so with the 64x48
    grid of a camera image the candidate set is the same in both readings: checked here on random configurations,
    together with the one observable property — integer bounds make the flag a no-op by construction."""
    d = np.stack([bits(0)])
    b2 = (-1.9, 638.1, -7.4, 487.2)
    rng = np.random.default_rng(0)
    for _ in range(2000):
        x = float(rng.uniform(0, 630)); y = float(rng.uniform(0, 470))
        u = x + float(rng.uniform(-30, 30)); v = y + float(rng.uniform(-30, 30))
        r = max(abs(x - u), abs(y - v)) + float(rng.choice([-0.05, 0.01, 0.3]))
        kk = np.array([kp(x, y, 1)], KP_DTYPE)
        pp = pts(1, [1], u=u, v=v, radius=r)
        a, _ = oracle.fuse_search(oracle.frame_view(kk, d, 640, 480, SCALE, bounds=b2), pp)
        c, _ = oracle.fuse_search(oracle.frame_view(kk, d, 640, 480, SCALE, bounds=b2, keyframe=True), pp)
        assert a[0] == c[0] == (0 if r > max(abs(np.float32(x) - np.float32(u)), abs(np.float32(y) - np.float32(v))) else -1)
