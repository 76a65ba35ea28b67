"""Python mirror of ``ORB_SLAM2::ORBmatcher`` (reference include/ORBmatcher.h:51-117) for the
routines on the hot path, on top of the C ABI.  Frames and map points are passed as plain
arrays (the "views" of include/orbfe.h); results come back as index arrays the caller writes
into its own Frame / MapPoint objects, exactly what the C++ adapter does.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import (KP_DTYPE, BowView, Camera, DeviceFrameView, FrameView, MapPointsView, FramePointsView, PyramidView,
                   TriSideView, WindowPointsView, GRID_CELLS, check)


class Frame:
    """The fields of ``ORB_SLAM2::Frame`` the matcher reads (include/Frame.h): undistorted
    keypoints, descriptors, optional ``mvuRight``, image bounds and scale factors."""

    def __init__(self, keypoints: np.ndarray, descriptors: np.ndarray, width: float, height: float,
                 scale_factors: np.ndarray, uright: np.ndarray | None = None,
                 min_x: float = 0.0, min_y: float = 0.0, keyframe: bool = False):
        # keyframe: the object stands for a KeyFrame, whose GetFeaturesInArea subtracts the int-truncated
        # mnMinX / mnMinY (include/KeyFrame.h:191-194) while the grid itself is the Frame's
        # width / height are mnMaxX / mnMaxY (the image size when the camera has no distortion)
        self.kps = np.ascontiguousarray(keypoints, KP_DTYPE)
        self.desc = np.ascontiguousarray(descriptors if descriptors is not None else np.zeros((0, 32)), np.uint8).reshape(-1, 32)
        assert len(self.kps) == len(self.desc)
        self.scale_factors = np.ascontiguousarray(scale_factors, np.float32)
        self.uright = None if uright is None else np.ascontiguousarray(uright, np.float32)
        self.min_x, self.max_x, self.min_y, self.max_y = float(min_x), float(width), float(min_y), float(height)
        self.keyframe = bool(keyframe)

    @property
    def N(self) -> int:
        return len(self.kps)

    def view(self) -> FrameView:
        return FrameView(self.N, self.kps.ctypes.data, self.desc.ctypes.data,
                         None if self.uright is None else self.uright.ctypes.data,
                         self.min_x, self.max_x, self.min_y, self.max_y,
                         self.scale_factors.ctypes.data, len(self.scale_factors), int(self.keyframe))


class BowSide:
    """One side of SearchByBoW: descriptors, keypoint angles, optional live-map-point flags and the
    DBoW2::FeatureVector as ``{node id: [feature indices]}`` (flattened in map order)."""

    def __init__(self, descriptors: np.ndarray, angles: np.ndarray, feat_vec: dict, valid: np.ndarray | None = None):
        self.desc = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32)
        self.angle = np.ascontiguousarray(angles, np.float32)
        assert len(self.desc) == len(self.angle)
        self.node_id = np.array(sorted(feat_vec), np.uint32)
        self.node_start = np.zeros(len(self.node_id) + 1, np.int32)
        for i, k in enumerate(self.node_id):
            self.node_start[i + 1] = self.node_start[i] + len(feat_vec[int(k)])
        self.indices = np.array([j for k in self.node_id for j in feat_vec[int(k)]], np.uint32)
        self.valid = None if valid is None else np.ascontiguousarray(valid, np.uint8)

    def view(self) -> BowView:
        return BowView(len(self.desc), self.desc.ctypes.data, self.angle.ctypes.data,
                       None if self.valid is None else self.valid.ctypes.data, len(self.node_id),
                       self.node_id.ctypes.data, self.node_start.ctypes.data, self.indices.ctypes.data)


def project_last_frame(Tcw_cur: np.ndarray, Tcw_last: np.ndarray, world_pos: np.ndarray, last_octave: np.ndarray,
                       has_point: np.ndarray, outlier: np.ndarray, *, fx: float, fy: float, cx: float, cy: float,
                       mbf: float, mb: float, bounds: tuple, scale_factors: np.ndarray, th: float, mono: bool) -> dict:
    """Host side of SearchByProjection(CurrentFrame, LastFrame, th, bMono), src/ORBmatcher.cc:1494-1553:
    pose projection of the last frame's map points and the per-point search parameters, in cv::Mat
    float32 semantics (3x3 * 3x1 products accumulate in double and round once, the translation is a
    float add).  Returns the arrays `ORBmatcher.SearchByProjectionLast` takes (without angle / desc /
    observations, which the caller adds) plus the forward / backward flags."""
    f32 = np.float32
    Tc, Tl = np.asarray(Tcw_cur, f32), np.asarray(Tcw_last, f32)
    X = np.asarray(world_pos, f32).reshape(-1, 3)
    mul = lambda A, b: (A.astype(np.float64) @ b.astype(np.float64)).astype(f32)      # cv::gemm, float data
    Rcw, tcw = Tc[:3, :3], Tc[:3, 3]
    twc = (-(Rcw.T.astype(np.float64) @ tcw.astype(np.float64))).astype(f32)             # :1497
    Rlw, tlw = Tl[:3, :3], Tl[:3, 3]
    tlc = (mul(Rlw, twc) + tlw).astype(f32)                                               # :1502
    forward = bool(tlc[2] > f32(mb)) and not mono                                         # :1505-1506
    backward = bool(-tlc[2] > f32(mb)) and not mono
    x3Dc = ((Rcw.astype(np.float64) @ X.T.astype(np.float64)).astype(f32).T + tcw).astype(f32)   # :1516
    xc, yc = x3Dc[:, 0], x3Dc[:, 1]
    with np.errstate(divide="ignore"):
        invzc = (1.0 / x3Dc[:, 2].astype(np.float64)).astype(f32)                         # :1520
    u = (f32(fx) * xc * invzc + f32(cx)).astype(f32)                                      # :1528-1529
    v = (f32(fy) * yc * invzc + f32(cy)).astype(f32)
    minx, maxx, miny, maxy = (f32(b) for b in bounds)
    valid = (np.asarray(has_point, bool) & ~np.asarray(outlier, bool) & ~(invzc < 0) &
             ~((u < minx) | (u > maxx)) & ~((v < miny) | (v > maxy)))                    # :1510-1538
    o = np.asarray(last_octave, np.int32)
    radius = (f32(th) * np.asarray(scale_factors, f32)[o]).astype(f32)                    # :1543
    if forward:
        lo, hi = o.copy(), np.full_like(o, -1)                                            # :1547-1548
    elif backward:
        lo, hi = np.zeros_like(o), o.copy()                                               # :1549-1550
    else:
        lo, hi = o - 1, o + 1                                                             # :1551-1552
    ur = (u - f32(mbf) * invzc).astype(f32)                                               # :1575
    return dict(valid=valid.astype(np.uint8), u=u, v=v, ur=ur, radius=radius, min_level=lo.astype(np.int32),
                max_level=hi.astype(np.int32), forward=forward, backward=backward)


class TriSide:
    """One key frame of ``SearchForTriangulation``: mvKeysUn, mDescriptors, "has a map point" flags, mvuRight
    (None = monocular) and mFeatVec ``{node id: [feature indices]}``, flattened for ``orbfe_tri_side``."""

    def __init__(self, keypoints: np.ndarray, descriptors: np.ndarray, has_mappoint: np.ndarray, feat_vec: dict, uright=None):
        assert keypoints.dtype == KP_DTYPE
        self.kps = np.ascontiguousarray(keypoints)
        self.desc = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32)
        self.has_mappoint = np.ascontiguousarray(has_mappoint, np.uint8)
        self.uright = None if uright is None else np.ascontiguousarray(uright, np.float32)
        self.node_id = np.array(sorted(feat_vec), np.uint32)
        self.node_start = np.zeros(len(self.node_id) + 1, np.int32)
        for i, k in enumerate(self.node_id):
            self.node_start[i + 1] = self.node_start[i] + len(feat_vec[int(k)])
        self.indices = np.array([j for k in self.node_id for j in feat_vec[int(k)]], np.uint32)

    def view(self) -> TriSideView:
        return TriSideView(len(self.kps), self.kps.ctypes.data, self.desc.ctypes.data, self.has_mappoint.ctypes.data,
                           None if self.uright is None else self.uright.ctypes.data, len(self.node_id),
                           self.node_id.ctypes.data, self.node_start.ctypes.data, self.indices.ctypes.data)


def project_map_points(Rcw: np.ndarray, tcw: np.ndarray, world_pos: np.ndarray, *, fx: float, fy: float, cx: float, cy: float,
                       bounds: tuple, scale_factors: np.ndarray, log_scale_factor: float, th: float,
                       max_distance: np.ndarray, min_distance: np.ndarray, usable: np.ndarray | None = None,
                       normals: np.ndarray | None = None, Ow: np.ndarray | None = None, bf: float = 0.0,
                       depth_gate: bool = True, invz_double: bool = True, frame_bounds: bool = False,
                       camera_norm: bool = False) -> dict:
    """Host side of the key-frame guided searches: the pose arithmetic and gates the reference runs per map
    point before it looks at keypoints, in cv::Mat float32 semantics (3x3 * 3x1 products accumulate in double
    and round once, cv::norm accumulates in double).  One routine, switched by the flags, covers
      SearchByProjection(F, pKF, ...)   src/ORBmatcher.cc:1663-1703  depth_gate=False, frame_bounds=True
      SearchByProjection(pKF, Scw, ..)  :361-407                     invz_double=False, normals given
      Fuse(pKF, vpMapPoints, th)        :955-996                     invz_double=False, normals, bf
      Fuse(pKF, Scw, ...)               :1127-1168                   normals given
      SearchBySim3 (one direction)      :1300-1331                   camera_norm=True (Rcw, tcw = the composed
                                                                     transform into the other key frame)
    max_distance / min_distance are GetMaxDistanceInvariance() / GetMinDistanceInvariance() (1.2 mfMaxDistance,
    0.8 mfMinDistance, src/MapPoint.cc:406-416); PredictScale (:418-433) divides mfMaxDistance = max_distance / 1.2f.
    Returns valid / u / v / ur / radius / level arrays (``usable`` folds isBad, already-found and NULL tests)."""
    f32, f64 = np.float32, np.float64
    R, t = np.asarray(Rcw, f32).reshape(3, 3), np.asarray(tcw, f32).reshape(3)
    X = np.asarray(world_pos, f32).reshape(-1, 3)
    n = len(X)
    Xc = ((R.astype(f64) @ X.T.astype(f64)).astype(f32).T + t).astype(f32)
    z = Xc[:, 2]
    valid = np.ones(n, bool) if usable is None else np.asarray(usable, bool).copy()
    if depth_gate:
        valid &= ~(z < 0)
    with np.errstate(divide="ignore", invalid="ignore"):
        invz = (1.0 / z.astype(f64)).astype(f32) if invz_double else (f32(1) / z).astype(f32)
        if frame_bounds:                                              # u = fx*xc*invzc + cx (:1675-1676)
            u = (f32(fx) * Xc[:, 0] * invz + f32(cx)).astype(f32); v = (f32(fy) * Xc[:, 1] * invz + f32(cy)).astype(f32)
        else:                                                         # x = X*invz; u = fx*x + cx
            u = (f32(fx) * (Xc[:, 0] * invz).astype(f32) + f32(cx)).astype(f32)
            v = (f32(fy) * (Xc[:, 1] * invz).astype(f32) + f32(cy)).astype(f32)
        ur = (u - f32(bf) * invz).astype(f32)
    minx, maxx, miny, maxy = (f32(b) for b in bounds)
    if frame_bounds:
        valid &= ~((u < minx) | (u > maxx) | (v < miny) | (v > maxy))
    else:                                                             # KeyFrame::IsInImage (KeyFrame.cc:653-656)
        valid &= (u >= minx) & (u < maxx) & (v >= miny) & (v < maxy)
    if camera_norm:
        PO = Xc
    else:
        if Ow is None:
            Ow = (-(R.T.astype(f64) @ t.astype(f64))).astype(f32)
        PO = (X - np.asarray(Ow, f32).reshape(3)).astype(f32)
    dist = np.sqrt((PO.astype(f64) ** 2).sum(axis=1)).astype(f32)
    maxd, mind = np.asarray(max_distance, f32), np.asarray(min_distance, f32)
    valid &= ~((dist < mind) | (dist > maxd))
    if normals is not None:                                           # PO.dot(Pn) < 0.5*dist
        dot = (PO.astype(f64) * np.asarray(normals, f32).reshape(-1, 3).astype(f64)).sum(axis=1)
        valid &= ~(dot < 0.5 * dist.astype(f64))
    with np.errstate(divide="ignore", invalid="ignore"):
        ratio = ((maxd / f32(1.2)).astype(f32) / dist).astype(f32)
        lvl = np.ceil(np.log(ratio).astype(f32) / f32(log_scale_factor))
    nl = len(scale_factors)
    lvl = np.clip(np.nan_to_num(lvl, nan=0.0, posinf=nl - 1, neginf=0.0), 0, nl - 1).astype(np.int32)
    radius = (f32(th) * np.asarray(scale_factors, f32)[lvl]).astype(f32)
    return dict(valid=valid.astype(np.uint8), u=np.nan_to_num(u), v=np.nan_to_num(v), ur=np.nan_to_num(ur), radius=radius, level=lvl)


class ORBmatcher:
    """``ORBmatcher(nnratio=0.6, checkOri=True)`` (src/ORBmatcher.cc:41)."""

    TH_HIGH, TH_LOW, HISTO_LENGTH = 100, 50, 30   # src/ORBmatcher.cc:37-39

    def __init__(self, nnratio: float = 0.6, checkOri: bool = True, *, device: int = 0):
        self._L = _lib.lib()
        self.mfNNratio, self.mbCheckOrientation = float(nnratio), bool(checkOri)
        self._h = C.c_void_p()
        check(self._L.orbfe_matcher_create(device, C.byref(self._h)), "orbfe_matcher_create")

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.orbfe_matcher_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def stream(self) -> int:
        return int(self._L.orbfe_matcher_stream(self._h) or 0)

    def sync(self) -> None:
        check(self._L.orbfe_matcher_sync(self._h))

    @staticmethod
    def DescriptorDistance(a: np.ndarray, b: np.ndarray) -> int:
        """src/ORBmatcher.cc:1828-1844 (host scalar, as in the reference)."""
        a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
        assert a.size == 32 and b.size == 32
        return _lib.lib().orbfe_hamming256(a.ctypes.data, b.ctypes.data)

    def knn2(self, query: np.ndarray, train: np.ndarray, th: int | None = None, ratio: float | None = None):
        """Brute-force best/second-best search + TH_LOW/ratio acceptance (src/ORBmatcher.cc:233-269).
        -> (idx int32[M] (-1 = rejected), d1 int32[M], d2 int32[M])"""
        q = np.ascontiguousarray(query, np.uint8).reshape(-1, 32)
        t = np.ascontiguousarray(train, np.uint8).reshape(-1, 32)
        th = self.TH_LOW if th is None else th
        ratio = self.mfNNratio if ratio is None else ratio
        m = len(q)
        idx, d1, d2 = (np.empty(m, np.int32) for _ in range(3))
        check(self._L.orbfe_knn2(self._h, q.ctypes.data, m, t.ctypes.data, len(t), th, ratio,
                                 idx.ctypes.data, d1.ctypes.data, d2.ctypes.data), "orbfe_knn2")
        return idx, d1, d2

    def knn2_device(self, q_ptr: int, m: int, t_ptr: int, n: int, idx_ptr: int, d1_ptr: int, d2_ptr: int,
                    th: int | None = None, ratio: float | None = None) -> None:
        th = self.TH_LOW if th is None else th
        ratio = self.mfNNratio if ratio is None else ratio
        check(self._L.orbfe_knn2_device(self._h, q_ptr, m, t_ptr, n, th, ratio, idx_ptr, d1_ptr, d2_ptr),
              "orbfe_knn2_device")

    def fetch(self, dst_ptr: int, src_dev_ptr: int, nbytes: int) -> None:
        """Asynchronous device-to-host copy on the matcher's stream (orbfe_matcher_fetch); sync() makes it visible."""
        check(self._L.orbfe_matcher_fetch(self._h, dst_ptr, src_dev_ptr, nbytes), "orbfe_matcher_fetch")

    KNN_ENGINES = {"auto": 0, "popc": 1, "mma": 2, "fp4": 3}

    def set_knn_engine(self, engine: str) -> None:
        """"popc": XOR + POPC on the integer pipe; "mma": tcgen05.mma kind::i8 on bit -> byte expanded rows; "fp4": tcgen05.mma
        kind::mxf4 on +-1 E2M1 rows; "auto" (default)."""
        check(self._L.orbfe_matcher_set_knn_engine(self._h, self.KNN_ENGINES[engine]), "orbfe_matcher_set_knn_engine")

    def knn2_sharded(self, comm, q_ptr: int, m: int, t_ptr: int, n: int, out_ptr: int,
                     th: int | None = None, ratio: float | None = None) -> None:
        """configs[4]: this rank's query shard on device pointers + one NCCL all-gather of 8-byte result rows
        (asynchronous on the matcher's stream; out_ptr: world * shard_rows(m, world) rows of KNN_RESULT_DTYPE)."""
        th = self.TH_LOW if th is None else th
        ratio = self.mfNNratio if ratio is None else ratio
        check(self._L.orbfe_knn2_sharded(self._h, comm._h, q_ptr, m, t_ptr, n, th, ratio, out_ptr), "orbfe_knn2_sharded")

    def SearchForInitialization(self, F1: Frame, F2: Frame, vbPrevMatched: np.ndarray, windowSize: int = 10):
        """src/ORBmatcher.cc:469-600.  vbPrevMatched float32 [N1,2] is updated in place.
        -> (nmatches, vnMatches12 int32[N1])"""
        assert vbPrevMatched.dtype == np.float32 and vbPrevMatched.flags.c_contiguous and vbPrevMatched.shape == (F1.N, 2)
        m12 = np.empty(max(F1.N, 1), np.int32)
        n = C.c_int()
        v1, v2 = F1.view(), F2.view()
        check(self._L.orbfe_search_init(self._h, C.byref(v1), C.byref(v2), vbPrevMatched.ctypes.data,
                                        m12.ctypes.data, windowSize, self.mfNNratio,
                                        int(self.mbCheckOrientation), C.byref(n)), "orbfe_search_init")
        return n.value, m12[:F1.N]

    def search_init_consecutive_device(self, kps_ptr: int, desc_ptr: int, counts_ptr: int, cap: int, nframes: int,
                                       width: float, height: float, m12_ptr: int, nm_ptr: int,
                                       windowSize: int = 100) -> None:
        """Batched SearchForInitialization over consecutive frames held in HBM (asynchronous)."""
        check(self._L.orbfe_search_init_consecutive_device(self._h, kps_ptr, desc_ptr, counts_ptr, cap, nframes,
                                                           0.0, float(width), 0.0, float(height), windowSize,
                                                           self.mfNNratio, int(self.mbCheckOrientation),
                                                           m12_ptr, nm_ptr), "orbfe_search_init_consecutive_device")

    def SearchByProjection(self, F: Frame, mappoints: dict, th: float = 3.0, occupied: np.ndarray | None = None):
        """SearchByProjection(Frame&, vector<MapPoint*>&, th) — src/ORBmatcher.cc:51-138.
        mappoints: dict of arrays track_in_view u8, proj_x/proj_y/proj_xr f32, level i32,
        view_cos f32, desc u8[n,32], observations i32.  occupied u8[N] in/out.
        -> (nmatches, mp_for_kp int32[N])"""
        spec = dict(track_in_view=np.uint8, proj_x=np.float32, proj_y=np.float32, proj_xr=np.float32,
                    level=np.int32, view_cos=np.float32, desc=np.uint8, observations=np.int32)
        keep = {k: np.ascontiguousarray(mappoints[k], t) for k, t in spec.items()}
        v = MapPointsView(len(keep["level"]), *[keep[k].ctypes.data for k in
                          ("track_in_view", "proj_x", "proj_y", "proj_xr", "level", "view_cos", "desc", "observations")])
        resident = isinstance(F, DeviceFrameView)
        N = F.n if resident else F.N
        if occupied is None:
            occupied = np.zeros(max(N, 1), np.uint8)
        out = np.full(max(N, 1), -1, np.int32)
        n = C.c_int()
        if resident:      # frame side already in HBM (ORBextractor.device_frame_view)
            check(self._L.orbfe_search_local_points_resident(self._h, C.byref(F), C.byref(v), th, self.mfNNratio,
                                                             occupied.ctypes.data, out.ctypes.data, C.byref(n)),
                  "orbfe_search_local_points_resident")
        else:
            fv = F.view()
            check(self._L.orbfe_search_local_points(self._h, C.byref(fv), C.byref(v), th, self.mfNNratio,
                                                    occupied.ctypes.data, out.ctypes.data, C.byref(n)),
                  "orbfe_search_local_points")
        return n.value, out[:N]

    FRAME_POINT_FIELDS = dict(valid=np.uint8, u=np.float32, v=np.float32, ur=np.float32, radius=np.float32,
                              min_level=np.int32, max_level=np.int32, angle=np.float32, desc=np.uint8,
                              observations=np.int32)

    def SearchByProjectionLast(self, F: Frame, points: dict, occupied: np.ndarray | None = None):
        """SearchByProjection(CurrentFrame, LastFrame, th, bMono) — src/ORBmatcher.cc:1484-1630, from the
        projected last-frame points (see `project_last_frame`).  occupied u8[N]: keypoints whose map
        point has observations.  -> (nmatches, match_for_kp int32[N]: >=0 last-frame index, -1
        untouched, -2 cleared by the rotation check)"""
        keep = {k: np.ascontiguousarray(points[k], t) for k, t in self.FRAME_POINT_FIELDS.items()}
        v = FramePointsView(len(keep["u"]), *[keep[k].ctypes.data for k in self.FRAME_POINT_FIELDS])
        resident = isinstance(F, DeviceFrameView)
        N = F.n if resident else F.N
        if occupied is None:
            occupied = np.zeros(max(N, 1), np.uint8)
        out = np.full(max(N, 1), -1, np.int32)
        n = C.c_int()
        if resident:
            check(self._L.orbfe_search_frame_points_resident(self._h, C.byref(F), C.byref(v), int(self.mbCheckOrientation),
                                                             occupied.ctypes.data, out.ctypes.data, C.byref(n)),
                  "orbfe_search_frame_points_resident")
        else:
            fv = F.view()
            check(self._L.orbfe_search_frame_points(self._h, C.byref(fv), C.byref(v), int(self.mbCheckOrientation),
                                                    occupied.ctypes.data, out.ctypes.data, C.byref(n)),
                  "orbfe_search_frame_points")
        return n.value, out[:N]

    WINDOW_POINT_FIELDS = dict(valid=np.uint8, u=np.float32, v=np.float32, ur=np.float32, radius=np.float32,
                               level=np.int32, desc=np.uint8)

    def _window_view(self, points: dict):
        n = len(points["u"])
        keep = {k: (np.ascontiguousarray(points[k], t) if points.get(k) is not None else np.zeros(max(n, 1), t))
                for k, t in self.WINDOW_POINT_FIELDS.items()}
        return WindowPointsView(n, *[keep[k].ctypes.data for k in self.WINDOW_POINT_FIELDS]), keep

    def _projected(self, F, points: dict, max_dist: int, check_ori: bool, occupied):
        keep = {k: np.ascontiguousarray(points[k], t) for k, t in self.FRAME_POINT_FIELDS.items()}
        v = FramePointsView(len(keep["u"]), *[keep[k].ctypes.data for k in self.FRAME_POINT_FIELDS])
        resident = isinstance(F, DeviceFrameView)
        N = F.n if resident else F.N
        occ = np.zeros(max(N, 1), np.uint8) if occupied is None else np.ascontiguousarray(occupied, np.uint8)
        out = np.full(max(N, 1), -1, np.int32)
        n = C.c_int()
        if resident:
            check(self._L.orbfe_search_projected_points_resident(self._h, C.byref(F), C.byref(v), int(max_dist), int(check_ori),
                                                                 occ.ctypes.data, out.ctypes.data, C.byref(n)),
                  "orbfe_search_projected_points_resident")
        else:
            fv = F.view()
            check(self._L.orbfe_search_projected_points(self._h, C.byref(fv), C.byref(v), int(max_dist), int(check_ori),
                                                        occ.ctypes.data, out.ctypes.data, C.byref(n)),
                  "orbfe_search_projected_points")
        return n.value, out[:N]

    def SearchByProjectionKeyFrame(self, F: Frame, points: dict, ORBdist: int, occupied: np.ndarray | None = None):
        """SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, sAlreadyFound, th, ORBdist) — src/ORBmatcher.cc:1641-1770.
        points: valid / u / v / radius / level (nPredictedLevel) / angle (pKF->mvKeysUn[i].angle) / desc per map point of
        pKF (see `project_map_points`); occupied u8[N]: CurrentFrame.mvpMapPoints[i] != NULL.
        -> (nmatches, match_for_kp int32[N]: >= 0 index i of pKF's map point, -1 untouched, -2 cleared by the rotation check)"""
        lvl = np.asarray(points["level"], np.int32)
        n = len(lvl)
        p = dict(points, min_level=lvl - 1, max_level=lvl + 1, observations=np.ones(n, np.int32),      # :1705, :1720
                 ur=np.zeros(n, np.float32))
        return self._projected(F, p, ORBdist, self.mbCheckOrientation, occupied)

    def SearchByProjectionSim3(self, KF: Frame, points: dict, matched: np.ndarray | None = None):
        """SearchByProjection(KeyFrame* pKF, Scw, vpPoints, vpMatched, th) — src/ORBmatcher.cc:335-451.  points as above
        without angle; matched u8[N]: vpMatched[idx] != NULL on entry.
        -> (nmatches, match_for_kp int32[N]: index into vpPoints newly written to vpMatched[idx], else -1)"""
        lvl = np.asarray(points["level"], np.int32)
        n = len(lvl)
        p = dict(points, min_level=lvl - 1, max_level=lvl, observations=np.ones(n, np.int32),           # :426, :421
                 ur=np.zeros(n, np.float32), angle=np.zeros(n, np.float32))
        return self._projected(KF, p, self.TH_LOW, False, matched)

    def FuseSearch(self, KF, points: dict, inv_level_sigma2: np.ndarray | None = None):
        """The search of Fuse(pKF, vpMapPoints, th) (src/ORBmatcher.cc:940-1090; pass pKF->mvInvLevelSigma2) and of
        Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) (:1100-1237; inv_level_sigma2 None).
        -> (best_idx int32[n]: keypoint to fuse with or -1, best_dist int32[n]); Replace / AddObservation stay with the caller."""
        v, keep = self._window_view(points)
        n = v.n
        bi = np.full(max(n, 1), -1, np.int32); bd = np.full(max(n, 1), 256, np.int32)
        sig = None if inv_level_sigma2 is None else np.ascontiguousarray(inv_level_sigma2, np.float32)
        sp = None if sig is None else sig.ctypes.data
        if isinstance(KF, DeviceFrameView):
            check(self._L.orbfe_search_window_best_resident(self._h, C.byref(KF), C.byref(v), sp, self.TH_LOW, bi.ctypes.data,
                                                            bd.ctypes.data), "orbfe_search_window_best_resident")
        else:
            fv = KF.view()
            check(self._L.orbfe_search_window_best(self._h, C.byref(fv), C.byref(v), sp, self.TH_LOW, bi.ctypes.data,
                                                   bd.ctypes.data), "orbfe_search_window_best")
        return bi[:n], bd[:n]

    def SearchBySim3(self, KF1: Frame, KF2: Frame, points1: dict, points2: dict):
        """SearchBySim3 — src/ORBmatcher.cc:1251-1477.  points1[i1]: KF1's map point i1 projected into KF2 (one entry per
        keypoint of KF1, valid = 0 where there is none / already matched / bad), points2 likewise.
        -> (nFound, matches12 int32[N1]: index in KF2 or -1)"""
        v1, k1 = self._window_view(points1); v2, k2 = self._window_view(points2)
        m12 = np.full(max(v1.n, 1), -1, np.int32)
        n = C.c_int()
        f1, f2 = KF1.view(), KF2.view()
        check(self._L.orbfe_search_by_sim3(self._h, C.byref(f1), C.byref(f2), C.byref(v1), C.byref(v2), self.TH_HIGH,
                                           m12.ctypes.data, C.byref(n)), "orbfe_search_by_sim3")
        return n.value, m12[:v1.n]

    def SearchForTriangulation(self, side1: TriSide, side2: TriSide, F12: np.ndarray, ex: float, ey: float,
                               scale_factors2: np.ndarray, level_sigma2_2: np.ndarray, bOnlyStereo: bool = False):
        """SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo) — src/ORBmatcher.cc:754-928; (ex, ey) is the
        epipole of :762-768.  -> (nmatches, matches12 int32[N1]); vMatchedPairs = [(i, m) for i, m in enumerate(matches12) if m >= 0]"""
        F = np.ascontiguousarray(F12, np.float32).reshape(9)
        sf = np.ascontiguousarray(scale_factors2, np.float32); sg = np.ascontiguousarray(level_sigma2_2, np.float32)
        n1 = len(side1.kps)
        m12 = np.full(max(n1, 1), -1, np.int32)
        n = C.c_int()
        v1, v2 = side1.view(), side2.view()
        check(self._L.orbfe_search_triangulation(self._h, C.byref(v1), C.byref(v2), F.ctypes.data, float(ex), float(ey),
                                                 sf.ctypes.data, sg.ctypes.data, len(sf), int(bOnlyStereo),
                                                 int(self.mbCheckOrientation), m12.ctypes.data, C.byref(n)),
              "orbfe_search_triangulation")
        return n.value, m12[:n1]

    def SearchByBoW(self, side1: BowSide, side2: BowSide, keyframe_pair: bool = False):
        """SearchByBoW(KeyFrame*, Frame&, ..) — src/ORBmatcher.cc:185-325 — or, with ``keyframe_pair``,
        SearchByBoW(KeyFrame*, KeyFrame*, ..) — :610-743 (bestDist1 < TH_LOW instead of <=).
        -> (nmatches, matches12 int32[n1], matches21 int32[n2]); partner index or -1."""
        n1, n2 = len(side1.desc), len(side2.desc)
        m12 = np.empty(max(n1, 1), np.int32); m21 = np.empty(max(n2, 1), np.int32)
        n = C.c_int()
        v1, v2 = side1.view(), side2.view()
        check(self._L.orbfe_search_bow(self._h, C.byref(v1), C.byref(v2), self.TH_LOW - 1 if keyframe_pair else self.TH_LOW,
                                       self.mfNNratio, int(self.mbCheckOrientation), m12.ctypes.data, m21.ctypes.data,
                                       C.byref(n)), "orbfe_search_bow")
        return n.value, m12[:n1], m21[:n2]

    def SearchByBoWBatch(self, sides1: list, sides2: list, keyframe_pair: bool = False):
        """``SearchByBoW`` for every (sides1[p], sides2[p]) in one call; a BowSide object that appears in
        several pairs is uploaded once (relocalisation: one frame against all candidate key frames).
        -> list of (nmatches, matches12, matches21)"""
        P = len(sides1)
        assert len(sides2) == P
        if P == 0:
            return []
        views = {}
        v1 = (BowView * P)(*[views.setdefault(id(s), s.view()) for s in sides1])
        v2 = (BowView * P)(*[views.setdefault(id(s), s.view()) for s in sides2])
        m12 = [np.empty(max(len(s.desc), 1), np.int32) for s in sides1]
        m21 = [np.empty(max(len(s.desc), 1), np.int32) for s in sides2]
        p12 = (C.c_void_p * P)(*[a.ctypes.data for a in m12]); p21 = (C.c_void_p * P)(*[a.ctypes.data for a in m21])
        nm = np.zeros(P, np.int32)
        check(self._L.orbfe_search_bow_batch(self._h, P, C.cast(v1, C.c_void_p), C.cast(v2, C.c_void_p),
                                             self.TH_LOW - 1 if keyframe_pair else self.TH_LOW, self.mfNNratio,
                                             int(self.mbCheckOrientation), C.cast(p12, C.c_void_p), C.cast(p21, C.c_void_p),
                                             nm.ctypes.data), "orbfe_search_bow_batch")
        return [(int(nm[p]), m12[p][:len(sides1[p].desc)], m21[p][:len(sides2[p].desc)]) for p in range(P)]

    def ComputeDistinctiveDescriptors(self, descriptors: np.ndarray, start: np.ndarray) -> np.ndarray:
        """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:266-334) for a batch of map points:
        point p owns ``descriptors[start[p]:start[p+1]]``.  -> int32[npoints]: index inside each set of
        the descriptor with the least median distance (-1 for an empty set)."""
        d = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32)
        st = np.ascontiguousarray(start, np.int32)
        npts = len(st) - 1
        best = np.empty(max(npts, 1), np.int32)
        check(self._L.orbfe_distinctive_descriptors(self._h, d.ctypes.data, st.ctypes.data, npts, best.ctypes.data),
              "orbfe_distinctive_descriptors")
        return best[:npts]

    # --- Frame constructor epilogue for callers without a batched extractor handle
    def ComputeImageBounds(self, camera, cols: int, rows: int) -> np.ndarray:
        """Frame::ComputeImageBounds (src/Frame.cc:481-512) -> float32 (mnMinX, mnMaxX, mnMinY, mnMaxY)."""
        cam = Camera(*(list(camera) + [0.0] * (9 - len(camera))))
        b = np.zeros(4, np.float32)
        check(self._L.orbfe_image_bounds(self._h, C.byref(cam), cols, rows, b.ctypes.data), "orbfe_image_bounds")
        return b

    def UndistortAndAssign(self, camera, bounds, keypoints: np.ndarray):
        """Frame::UndistortKeyPoints + AssignFeaturesToGrid (src/Frame.cc:443-475, :246-267).
        -> (mvKeysUn KP_DTYPE[n], cell_end int32[3072], items uint16[k]); mGrid[ix][iy] =
        items[cell_end[c-1]:cell_end[c]], c = ix*48 + iy."""
        cam = Camera(*(list(camera) + [0.0] * (9 - len(camera))))
        b = np.ascontiguousarray(bounds, np.float32)
        kps = np.ascontiguousarray(keypoints, KP_DTYPE)
        n = len(kps)
        un = np.empty(max(n, 1), KP_DTYPE); ce = np.empty(GRID_CELLS, np.int32); it = np.empty(max(n, 1), np.uint16)
        k = C.c_int()
        check(self._L.orbfe_frame_assign(self._h, C.byref(cam), b.ctypes.data, kps.ctypes.data, n, un.ctypes.data,
                                         ce.ctypes.data, it.ctypes.data, C.byref(k)), "orbfe_frame_assign")
        return un[:n], ce, it[:k.value]

    def ComputeStereoMatchesBatch(self, extractor, npairs: int, mbf: float, mb: float):
        """Frame::ComputeStereoMatches (src/Frame.cc:512-686) for the ``npairs`` stereo pairs of the extractor's last batch
        (frames 2p / 2p+1 = left / right of pair p), read where they lie in device memory.
        -> (mvuRight float32[npairs, cap], mvDepth float32[npairs, cap]); cap = extractor.max_keypoints"""
        kp, ds, cn, cap = extractor.device_results()
        pv = extractor.device_pyramid()
        sf = np.ascontiguousarray(extractor.GetScaleFactors(), np.float32)
        isf = np.ascontiguousarray(extractor.GetInverseScaleFactors(), np.float32)
        ur = np.empty((max(npairs, 1), cap), np.float32); dp = np.empty((max(npairs, 1), cap), np.float32)
        check(self._L.orbfe_stereo_match_batch_device(self._h, C.byref(pv), kp, ds, cn, cap, npairs, sf.ctypes.data, isf.ctypes.data,
                                                      len(sf), mbf, mb, ur.ctypes.data, dp.ctypes.data),
              "orbfe_stereo_match_batch_device")
        return ur[:npairs], dp[:npairs]

    def ComputeStereoMatches(self, left: Frame, right: Frame, pyr_left: list, pyr_right: list,
                             inv_scale_factors: np.ndarray, mbf: float, mb: float):
        """Frame::ComputeStereoMatches — src/Frame.cc:512-686. -> (mvuRight f32[N], mvDepth f32[N])"""
        def pv(levels):
            lv = [np.ascontiguousarray(l, np.uint8) for l in levels]
            ptrs = (C.c_void_p * len(lv))(*[l.ctypes.data for l in lv])
            w = np.array([l.shape[1] for l in lv], np.int32); h = np.array([l.shape[0] for l in lv], np.int32)
            st = np.array([l.strides[0] for l in lv], np.int64)
            return PyramidView(len(lv), C.cast(ptrs, C.c_void_p), w.ctypes.data, h.ctypes.data, st.ctypes.data), (lv, ptrs, w, h, st)
        pl, k1 = pv(pyr_left); pr, k2 = pv(pyr_right)
        inv = np.ascontiguousarray(inv_scale_factors, np.float32)
        ur = np.empty(max(left.N, 1), np.float32); dp = np.empty(max(left.N, 1), np.float32)
        lv, rv = left.view(), right.view()
        check(self._L.orbfe_stereo_match(self._h, C.byref(lv), C.byref(rv), C.byref(pl), C.byref(pr),
                                         inv.ctypes.data, mbf, mb, ur.ctypes.data, dp.ctypes.data),
              "orbfe_stereo_match")
        return ur[:left.N], dp[:left.N]
