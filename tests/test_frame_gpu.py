"""GPU parity of the Frame epilogue (K11: Frame::UndistortKeyPoints + ComputeImageBounds +
AssignFeaturesToGrid) through the C ABI: undistorted positions bit-exact against the committed
cv2 vectors and the oracle, the grid identical to the oracle's, and the guided searches on a
device-resident frame identical to the same searches on uploaded arrays."""
import os

import numpy as np
import pytest

from orb_slam2_annotate_b200 import synth

pytestmark = pytest.mark.gpu

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "undistort_cv2.npz"))
CAMERAS = ("tum1", "tum2", "euroc", "wide")


def _camera(name):
    return list(G[name + "_cam"]) + list(G[name + "_dist"])


def _grid_lists(cell_end, items):
    start = np.concatenate([[0], cell_end[:-1]])
    return [list(items[s:e]) for s, e in zip(start, cell_end)]


def _oracle_grid(oracle, kps_un, bounds):
    f = oracle.frame_view(kps_un, np.zeros((len(kps_un), 32), np.uint8), 0, 0, np.ones(8, np.float32), bounds=bounds)
    cs, items = oracle.assign_grid(f)
    return [list(items[cs[c]:cs[c + 1]]) for c in range(64 * 48)]


@pytest.mark.parametrize("name", CAMERAS)
def test_undistort_and_grid_golden_points(orbfe, oracle, name):
    cols, rows = (int(v) for v in G[name + "_size"])
    mt = orbfe.ORBmatcher()
    bounds = mt.ComputeImageBounds(_camera(name), cols, rows)
    assert bounds.tobytes() == oracle.image_bounds(G[name + "_cam"], G[name + "_dist"], cols, rows).tobytes()
    pts = G[name + "_pts"]
    k = np.zeros(len(pts), orbfe.KP_DTYPE)
    k["x"], k["y"] = pts[:, 0], pts[:, 1]
    k["octave"] = np.arange(len(pts)) % 8; k["angle"] = np.arange(len(pts)) % 360; k["response"] = 7; k["size"] = 31
    un, cell_end, items = mt.UndistortAndAssign(_camera(name), bounds, k)
    got = np.stack([un["x"], un["y"]], 1)
    assert got.tobytes() == G[name + "_und"].tobytes()                       # cv2 itself
    for fld in ("size", "angle", "response", "octave", "class_id"):          # kp = mvKeys[i] otherwise (:469)
        assert np.array_equal(un[fld], k[fld])
    assert _grid_lists(cell_end, items) == _oracle_grid(oracle, un, bounds)
    assert 0 < len(items) <= len(pts)
    if name != "wide":
        assert len(items) < len(pts)      # the out-of-image samples fall off the grid (wide keeps the raw image box)


def test_frame_assign_edge_cases(orbfe, oracle):
    mt = orbfe.ORBmatcher()
    cam = _camera("tum1")
    b = mt.ComputeImageBounds(cam, 640, 480)
    un, cell_end, items = mt.UndistortAndAssign(cam, b, np.zeros(0, orbfe.KP_DTYPE))
    assert len(un) == 0 and len(items) == 0 and not cell_end.any()
    # k1 == 0: positions are copied, bounds are the image (Frame.cc:445-450, :505-510)
    cam0 = list(cam); cam0[4] = 0.0
    assert np.array_equal(mt.ComputeImageBounds(cam0, 752, 480), np.array([0, 752, 0, 480], np.float32))
    k = np.zeros(5000, orbfe.KP_DTYPE)
    rng = np.random.default_rng(0)
    k["x"] = rng.uniform(-10, 760, 5000); k["y"] = rng.uniform(-10, 490, 5000)
    k["x"][:64] = 20.0; k["y"][:64] = 20.0                 # one crowded cell: push_back order must survive
    un, cell_end, items = mt.UndistortAndAssign(cam0, (0, 752, 0, 480), k)
    assert un.tobytes() == k.tobytes()
    assert _grid_lists(cell_end, items) == _oracle_grid(oracle, k, (0, 752, 0, 480))
    with pytest.raises(orbfe.OrbfeError):
        mt.UndistortAndAssign(cam, (0, 0, 0, 480), k)


@pytest.mark.parametrize("name,B", [("tum1", 5), ("euroc", 3), (None, 2)])
def test_extractor_epilogue(orbfe, oracle, name, B):
    W, H = (752, 480) if name == "euroc" else (640, 480)
    imgs, _ = synth.sequence(B, W, H, seed=4)
    ex = orbfe.ORBextractor(1000, 1.2, 8, 20, 7)
    cam = _camera(name) if name else [500.0, 500.0, W / 2, H / 2, 0.0, 0.1, 0.0, 0.0, 0.0]   # k1 = 0: no undistortion
    bounds = ex.set_camera(cam, W, H)
    want_b = oracle.image_bounds(np.array(cam[:4], np.float32), np.array(cam[4:], np.float32), W, H)
    assert bounds.tobytes() == want_b.tobytes()
    kps, desc, counts = ex.extract_batch(imgs)
    un, cell_end, items = ex.fetch_frames()
    assert ex.stage_times_ms is not None
    for b in range(B):
        n = int(counts[b])
        assert n > 500
        xy = np.stack([kps[b, :n]["x"], kps[b, :n]["y"]], 1)
        want = oracle.undistort_points(xy, np.array(cam[:4], np.float32), np.array(cam[4:], np.float32))
        got = np.stack([un[b, :n]["x"], un[b, :n]["y"]], 1)
        assert got.tobytes() == want.tobytes()
        if name is None:
            assert un[b, :n].tobytes() == kps[b, :n].tobytes()
        else:
            assert (got != xy).any()
        for fld in ("size", "angle", "response", "octave", "class_id"):
            assert np.array_equal(un[b, :n][fld], kps[b, :n][fld])
        k_in = int(cell_end[b, -1])
        assert _grid_lists(cell_end[b], items[b, :k_in]) == _oracle_grid(oracle, un[b, :n], bounds)
    # switching the camera off again stops the epilogue but keeps extraction identical
    ex.set_camera(None, 0, 0)
    kps2, desc2, counts2 = ex.extract_batch(imgs)
    assert np.array_equal(counts, counts2) and all(kps[b, :counts[b]].tobytes() == kps2[b, :counts[b]].tobytes() for b in range(B))
    with pytest.raises(orbfe.OrbfeError):
        ex.device_frames()


def _mappoints(k, d, n, seed):
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


def test_guided_searches_on_resident_frames(orbfe, oracle):
    """SearchByProjection(F, MapPoints) and SearchByProjection(Current, Last) against frames that never
    left the device: same answers as the oracle fed with the fetched mvKeysUn."""
    B, W, H = 3, 640, 480
    imgs, _ = synth.sequence(B, W, H, seed=6)
    ex = orbfe.ORBextractor(1500, 1.2, 8, 20, 7)
    cam = _camera("tum1")
    bounds = ex.set_camera(cam, W, H)
    kps, desc, counts = ex.extract_batch(imgs)
    un, _, _ = ex.fetch_frames()
    scale = ex.GetScaleFactors()
    mt = orbfe.ORBmatcher(0.8, True)
    rng = np.random.default_rng(5)
    for b in (0, 2):
        n = int(counts[b])
        k, d = un[b, :n], desc[b, :n]
        dv = ex.device_frame_view(b, n)
        # local map points
        mps = _mappoints(k, d, 6000, 11 + b)
        occ = (rng.random(n) < 0.05).astype(np.uint8)
        occ_g, occ_h, occ_o = occ.copy(), occ.copy(), occ.copy()
        ng, mg = mt.SearchByProjection(dv, mps, 3.0, occ_g)
        Fh = orbfe.Frame(k, d, bounds[1], bounds[3], scale, min_x=bounds[0], min_y=bounds[2])
        nh, mh = mt.SearchByProjection(Fh, mps, 3.0, occ_h)
        no, mo = oracle.search_by_projection(oracle.frame_view(k, d, W, H, scale, bounds=bounds), mps, 3.0, 0.8, occ_o)
        assert ng == no == nh and np.array_equal(mg, mo) and np.array_equal(mh, mo) and np.array_equal(occ_g, occ_o)
        assert ng > 100
        # motion-model search: points of another frame of the batch projected near their own positions
        a = (b + 1) % B
        na = int(counts[a])
        ka, da = un[a, :na], desc[a, :na]
        pts = dict(valid=(rng.random(na) < 0.9).astype(np.uint8), u=(ka["x"] + rng.uniform(-2, 2, na)).astype(np.float32),
                   v=(ka["y"] + rng.uniform(-2, 2, na)).astype(np.float32), ur=np.zeros(na, np.float32),
                   radius=(15.0 * scale[ka["octave"]]).astype(np.float32), min_level=(ka["octave"] - 1).astype(np.int32),
                   max_level=(ka["octave"] + 1).astype(np.int32), angle=ka["angle"].astype(np.float32), desc=da,
                   observations=rng.integers(0, 3, na).astype(np.int32))
        ng, mg = mt.SearchByProjectionLast(dv, pts, occ.copy())
        no, mo = oracle.search_by_projection_frame(oracle.frame_view(k, d, W, H, scale, bounds=bounds), pts, True, occ)
        assert ng == no and np.array_equal(mg, mo)
        assert ng > 20
        # the key-frame overloads on the resident frame: SearchByProjection(F, KF) and the Fuse search (TUM1 is a
        # distorted camera, so the bounds are fractional and the KeyFrame reading truncates them)
        kp = dict(valid=pts["valid"], u=pts["u"], v=pts["v"], radius=(10.0 * scale[ka["octave"]]).astype(np.float32),
                  level=ka["octave"].astype(np.int32), angle=pts["angle"], desc=da)
        ng, mg = mt.SearchByProjectionKeyFrame(dv, kp, 80, occ.copy())
        no, mo = oracle.search_by_projection_keyframe(oracle.frame_view(k, d, W, H, scale, bounds=bounds), kp, 80, True, occ)
        assert ng == no and np.array_equal(mg, mo) and ng > 20
        isig = (np.float32(1.0) / (scale * scale)).astype(np.float32)
        for keyframe in (0, 1):
            dv.keyframe_bounds = keyframe
            for sig in (isig, None):             # with Fuse's reprojection gate (few survive: the points come from another frame) / without
                bi, bd = mt.FuseSearch(dv, kp, sig)
                oi, od = oracle.fuse_search(oracle.frame_view(k, d, W, H, scale, bounds=bounds, keyframe=bool(keyframe)), kp, sig)
                assert np.array_equal(bi, oi) and np.array_equal(bd, od) and (bi >= 0).sum() > (0 if sig is not None else 20)
        dv.keyframe_bounds = 0
