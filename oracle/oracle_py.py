"""ctypes bindings for the CPU oracle (liborb_oracle.so) and the compiled reference TU
(oracle/_ref/liborbref*.so).  TEST INFRASTRUCTURE ONLY: imported by tests/, by
__graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference legs, never by the
product package."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28

_u8p = C.POINTER(C.c_uint8)
_ip = C.POINTER(C.c_int)
_fp = C.POINTER(C.c_float)


def _p(a, t=_u8p):
    return a.ctypes.data_as(t)


def build(force: bool = False) -> None:
    """Compile liborb_oracle.so (and oracle/_ref when /root/reference is mounted)."""
    if force or not os.path.exists(os.path.join(HERE, "liborb_oracle.so")) or os.path.exists("/root/reference/src/ORBextractor.cc"):
        subprocess.run(["make", "-C", HERE, "all"], check=True, stdout=subprocess.DEVNULL)


class FrameView(C.Structure):
    _fields_ = [("n", C.c_int), ("kps", C.c_void_p), ("desc", C.c_void_p), ("uright", C.c_void_p),
                ("min_x", C.c_float), ("max_x", C.c_float), ("min_y", C.c_float), ("max_y", C.c_float),
                ("scale_factors", C.c_void_p), ("nlevels", C.c_int), ("keyframe_bounds", C.c_int)]


class MapPointsView(C.Structure):
    _fields_ = [("n", C.c_int), ("track_in_view", C.c_void_p), ("proj_x", C.c_void_p),
                ("proj_y", C.c_void_p), ("proj_xr", C.c_void_p), ("level", C.c_void_p),
                ("view_cos", C.c_void_p), ("desc", C.c_void_p), ("observations", C.c_void_p)]


class FramePointsView(C.Structure):
    _fields_ = [("n", C.c_int), ("valid", C.c_void_p), ("u", C.c_void_p), ("v", C.c_void_p),
                ("ur", C.c_void_p), ("radius", C.c_void_p), ("min_level", C.c_void_p),
                ("max_level", C.c_void_p), ("angle", C.c_void_p), ("desc", C.c_void_p),
                ("observations", C.c_void_p)]


class BowView(C.Structure):
    _fields_ = [("n", C.c_int), ("desc", C.c_void_p), ("angle", C.c_void_p), ("valid", C.c_void_p),
                ("nnodes", C.c_int), ("node_id", C.c_void_p), ("node_start", C.c_void_p), ("indices", C.c_void_p)]


class ProjPointsView(C.Structure):
    _fields_ = [("n", C.c_int), ("valid", C.c_void_p), ("u", C.c_void_p), ("v", C.c_void_p),
                ("ur", C.c_void_p), ("radius", C.c_void_p), ("level", C.c_void_p), ("angle", C.c_void_p),
                ("desc", C.c_void_p)]


class TriSide(C.Structure):
    _fields_ = [("n", C.c_int), ("kps", C.c_void_p), ("desc", C.c_void_p), ("has_mappoint", C.c_void_p),
                ("uright", C.c_void_p), ("nnodes", C.c_int), ("node_id", C.c_void_p), ("node_start", C.c_void_p),
                ("indices", C.c_void_p)]


class PyramidView(C.Structure):
    _fields_ = [("nlevels", C.c_int), ("data", C.c_void_p), ("w", C.c_void_p), ("h", C.c_void_p),
                ("stride", C.c_void_p)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(HERE, "liborb_oracle.so")
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        L.orc_fast_atan2.restype = C.c_float
        L.orc_fast_atan2.argtypes = [C.c_float, C.c_float]
        L.orc_cv_round_f.argtypes = [C.c_float]
        L.orc_cv_round_d.argtypes = [C.c_double]
        L.orc_extractor_create.restype = C.c_void_p
        L.orc_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.orc_extractor_destroy.argtypes = [C.c_void_p]
        L.orc_extractor_tables.argtypes = [C.c_void_p] + [C.c_void_p] * 6
        L.orc_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_ssize_t, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_level_size.argtypes = [C.c_void_p, C.c_int, _ip, _ip]
        L.orc_level_image.restype = C.c_void_p
        L.orc_level_image.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_ssize_t)]
        L.orc_level_blurred.restype = C.c_void_p
        L.orc_level_blurred.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_ssize_t)]
        L.orc_level_raw.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p)]
        L.orc_resize_linear_u8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_ssize_t, C.c_void_p, C.c_int, C.c_int, C.c_ssize_t]
        L.orc_copy_make_border_reflect101.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_ssize_t, C.c_void_p, C.c_ssize_t, C.c_int]
        L.orc_fast9_16.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_ssize_t, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_gaussian_blur_7x7_s2.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_ssize_t, C.c_void_p, C.c_ssize_t]
        L.orc_distribute_octtree.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_descriptor_distance.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_assign_grid.argtypes = [C.POINTER(FrameView), C.c_void_p, C.c_void_p]
        L.orc_features_in_area.argtypes = [C.POINTER(FrameView), C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_three_maxima.argtypes = [C.c_void_p, C.c_int, _ip, _ip, _ip]
        L.orc_search_for_initialization.argtypes = [C.POINTER(FrameView), C.POINTER(FrameView), C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_int]
        L.orc_search_by_projection_frame.argtypes = [C.POINTER(FrameView), C.POINTER(FramePointsView), C.c_int, C.c_void_p, C.c_void_p]
        L.orc_search_by_projection_frame.restype = C.c_int
        L.orc_search_by_projection.argtypes = [C.POINTER(FrameView), C.POINTER(MapPointsView), C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        L.orc_compute_stereo_matches.argtypes = [C.POINTER(FrameView), C.POINTER(FrameView), C.POINTER(PyramidView), C.POINTER(PyramidView), C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        L.orc_search_by_bow.argtypes = [C.POINTER(BowView), C.POINTER(BowView), C.c_int, C.c_float, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_sbp_frame_keyframe.argtypes = [C.POINTER(FrameView), C.POINTER(ProjPointsView), C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_sbp_keyframe_sim3.argtypes = [C.POINTER(FrameView), C.POINTER(ProjPointsView), C.c_void_p, C.c_void_p]
        L.orc_fuse_search.argtypes = [C.POINTER(FrameView), C.POINTER(ProjPointsView), C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_fuse_search.restype = None
        L.orc_search_by_sim3.argtypes = [C.POINTER(FrameView), C.POINTER(FrameView), C.POINTER(ProjPointsView), C.POINTER(ProjPointsView), C.c_void_p]
        L.orc_search_for_triangulation.argtypes = [C.POINTER(TriSide), C.POINTER(TriSide), C.c_void_p, C.c_float, C.c_float,
                                                   C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.orc_distinctive_descriptors.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orc_undistort_points.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_image_bounds.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.orc_knn2.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


# ----------------------------------------------------------------------------- primitives
def resize_linear(src: np.ndarray, dw: int, dh: int) -> np.ndarray:
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.empty((dh, dw), np.uint8)
    lib().orc_resize_linear_u8(src.ctypes.data, src.shape[1], src.shape[0], src.strides[0], dst.ctypes.data, dw, dh, dw)
    return dst


def copy_make_border(src: np.ndarray, b: int) -> np.ndarray:
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.empty((src.shape[0] + 2 * b, src.shape[1] + 2 * b), np.uint8)
    lib().orc_copy_make_border_reflect101(src.ctypes.data, src.shape[1], src.shape[0], src.strides[0], dst.ctypes.data, dst.strides[0], b)
    return dst


def fast(img: np.ndarray, threshold: int, nonmax: bool = True) -> np.ndarray:
    """-> int32 [n,3] (x,y,score) row-major."""
    assert img.dtype == np.uint8 and img.strides[1] == 1
    cap = max(1, img.shape[0] * img.shape[1])
    out = np.empty((cap, 3), np.int32)
    n = lib().orc_fast9_16(img.ctypes.data, img.shape[1], img.shape[0], img.strides[0], threshold, int(nonmax), out.ctypes.data, cap)
    return out[:n].copy()


def gaussian_blur(src: np.ndarray) -> np.ndarray:
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.empty_like(src)
    lib().orc_gaussian_blur_7x7_s2(src.ctypes.data, src.shape[1], src.shape[0], src.strides[0], dst.ctypes.data, dst.strides[0])
    return dst


def fast_atan2(y: float, x: float) -> float:
    return lib().orc_fast_atan2(y, x)


def distribute_octtree(xys: np.ndarray, minX, maxX, minY, maxY, N) -> np.ndarray:
    xys = np.ascontiguousarray(xys, np.int32)
    n = xys.shape[0]
    out = np.empty(n + 8, np.int32)
    c = lib().orc_distribute_octtree(xys.ctypes.data, n, minX, maxX, minY, maxY, N, out.ctypes.data, n + 8)
    return out[:c].copy()


def descriptor_distance(a: np.ndarray, b: np.ndarray) -> int:
    a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
    return lib().orc_descriptor_distance(a.ctypes.data, b.ctypes.data)


# ----------------------------------------------------------------------------- extractor
class OracleExtractor:
    """The repo's own CPU restatement of ORBextractor (oracle/orc_extractor.cpp)."""

    def __init__(self, nfeatures=1000, scale=1.2, nlevels=8, ini_th=20, min_th=7):
        self.nlevels = nlevels
        self.nfeatures = nfeatures
        self._h = lib().orc_extractor_create(nfeatures, scale, nlevels, ini_th, min_th)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_extractor_destroy(self._h); self._h = None

    def tables(self):
        n = self.nlevels
        s, i, g, ig = (np.empty(n, np.float32) for _ in range(4))
        f = np.empty(n, np.int32); u = np.empty(16, np.int32)
        lib().orc_extractor_tables(self._h, s.ctypes.data, i.ctypes.data, g.ctypes.data, ig.ctypes.data, f.ctypes.data, u.ctypes.data)
        return dict(scale=s, inv_scale=i, sigma2=g, inv_sigma2=ig, features_per_level=f, umax=u)

    def __call__(self, img: np.ndarray):
        assert img.dtype == np.uint8 and img.ndim == 2 and img.strides[1] == 1
        cap = 4 * self.nfeatures + 4096
        kps = np.empty(cap, KP_DTYPE); desc = np.empty((cap, 32), np.uint8)
        n = lib().orc_extract(self._h, img.ctypes.data, img.shape[1], img.shape[0], img.strides[0], kps.ctypes.data, desc.ctypes.data, cap)
        assert n >= 0
        return kps[:n].copy(), desc[:n].copy()

    def level(self, l: int, blurred=False):
        w = C.c_int(); h = C.c_int(); st = C.c_ssize_t()
        lib().orc_level_size(self._h, l, C.byref(w), C.byref(h))
        p = (lib().orc_level_blurred if blurred else lib().orc_level_image)(self._h, l, C.byref(st))
        if not p:
            return None
        buf = (C.c_uint8 * (w.value * h.value)).from_address(p)
        return np.frombuffer(buf, np.uint8).reshape(h.value, w.value).copy()

    def raw(self, l: int) -> np.ndarray:
        p = C.c_void_p()
        n = lib().orc_level_raw(self._h, l, C.byref(p))
        if n == 0:
            return np.zeros((0, 3), np.int32)
        buf = (C.c_int * (3 * n)).from_address(p.value)
        return np.frombuffer(buf, np.int32).reshape(n, 3).copy()


_ref_libs = {}


def ref_lib(variant: str = ""):
    """oracle/_ref/liborbref{variant}.so — the unmodified reference TU. None if not built."""
    if variant not in _ref_libs:
        path = os.path.join(HERE, "_ref", f"liborbref{variant}.so")
        if not os.path.exists(path):
            _ref_libs[variant] = None
        else:
            L = C.CDLL(path)
            L.ref_extractor_create.restype = C.c_void_p
            L.ref_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
            L.ref_extractor_destroy.argtypes = [C.c_void_p]
            L.ref_tables.argtypes = [C.c_void_p] * 5
            L.ref_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_ssize_t, C.c_void_p, C.c_void_p, C.c_int]
            L.ref_level.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), _ip, _ip, C.POINTER(C.c_ssize_t)]
            _ref_libs[variant] = L
    return _ref_libs[variant]


class RefExtractor:
    """The reference's own ORB_SLAM2::ORBextractor (compiled against the cv-shim)."""

    def __init__(self, nfeatures=1000, scale=1.2, nlevels=8, ini_th=20, min_th=7, variant=""):
        self._L = ref_lib(variant)
        if self._L is None:
            raise FileNotFoundError("oracle/_ref/liborbref%s.so not built" % variant)
        self.nlevels = nlevels; self.nfeatures = nfeatures
        self._h = self._L.ref_extractor_create(nfeatures, scale, nlevels, ini_th, min_th)

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.ref_extractor_destroy(self._h); self._h = None

    def tables(self):
        n = self.nlevels
        s, i, g, ig = (np.empty(n, np.float32) for _ in range(4))
        self._L.ref_tables(self._h, s.ctypes.data, i.ctypes.data, g.ctypes.data, ig.ctypes.data)
        return dict(scale=s, inv_scale=i, sigma2=g, inv_sigma2=ig)

    def __call__(self, img: np.ndarray):
        assert img.dtype == np.uint8 and img.ndim == 2 and img.strides[1] == 1
        cap = 4 * self.nfeatures + 4096
        kps = np.empty(cap, KP_DTYPE); desc = np.empty((cap, 32), np.uint8)
        n = self._L.ref_extract(self._h, img.ctypes.data, img.shape[1], img.shape[0], img.strides[0], kps.ctypes.data, desc.ctypes.data, cap)
        assert n >= 0
        return kps[:n].copy(), desc[:n].copy()

    def level(self, l: int) -> np.ndarray:
        p = C.c_void_p(); w = C.c_int(); h = C.c_int(); st = C.c_ssize_t()
        self._L.ref_level(self._h, l, C.byref(p), C.byref(w), C.byref(h), C.byref(st))
        buf = (C.c_uint8 * (st.value * h.value)).from_address(p.value)
        return np.frombuffer(buf, np.uint8).reshape(h.value, st.value)[:, :w.value].copy()


# ----------------------------------------------------------------------------- matcher
def frame_view(kps: np.ndarray, desc: np.ndarray, width: float, height: float, scale_factors: np.ndarray, uright=None,
               bounds=None, keyframe=False):
    """Returns (FrameView, keepalive list).  bounds = (mnMinX, mnMaxX, mnMinY, mnMaxY) overrides (0, width, 0, height)."""
    kps = np.ascontiguousarray(kps); desc = np.ascontiguousarray(desc, np.uint8)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    keep = [kps, desc, sf]
    v = FrameView(len(kps), kps.ctypes.data, desc.ctypes.data, None, 0.0, float(width), 0.0, float(height), sf.ctypes.data, len(sf))
    if uright is not None:
        ur = np.ascontiguousarray(uright, np.float32); keep.append(ur); v.uright = ur.ctypes.data
    if bounds is not None:
        v.min_x, v.max_x, v.min_y, v.max_y = (float(b) for b in bounds)
    v.keyframe_bounds = int(keyframe)
    return v, keep


def search_for_initialization(f1, f2, prev_matched: np.ndarray, window=100, nnratio=0.9, check_ori=True):
    """f1,f2 = (FrameView, keep). prev_matched float32[n1,2] is updated in place.
    -> (nmatches, matches12 int32[n1])"""
    n1 = f1[0].n
    m12 = np.empty(max(n1, 1), np.int32)
    assert prev_matched.dtype == np.float32 and prev_matched.flags.c_contiguous
    nm = lib().orc_search_for_initialization(C.byref(f1[0]), C.byref(f2[0]), prev_matched.ctypes.data, m12.ctypes.data, window, nnratio, int(check_ori))
    return nm, m12[:n1]


def search_by_projection(f, mps: dict, th: float, nnratio: float, occupied: np.ndarray):
    """mps: dict of arrays (track_in_view u8, proj_x/proj_y/proj_xr f32, level i32, view_cos f32,
    desc u8[n,32], observations i32). occupied u8[n] in/out. -> (nmatches, mp_for_kp int32[n])"""
    arrs = dict(track_in_view=np.uint8, proj_x=np.float32, proj_y=np.float32, proj_xr=np.float32,
                level=np.int32, view_cos=np.float32, desc=np.uint8, observations=np.int32)
    keep = {k: np.ascontiguousarray(mps[k], t) for k, t in arrs.items()}
    v = MapPointsView(len(keep["level"]), *[keep[k].ctypes.data for k in
                      ("track_in_view", "proj_x", "proj_y", "proj_xr", "level", "view_cos", "desc", "observations")])
    out = np.full(max(f[0].n, 1), -1, np.int32)
    nm = lib().orc_search_by_projection(C.byref(f[0]), C.byref(v), th, nnratio, occupied.ctypes.data, out.ctypes.data)
    return nm, out[:f[0].n]


def search_by_projection_frame(f, points: dict, check_ori: bool, occupied: np.ndarray):
    """points: dict of arrays valid u8, u/v/ur/radius f32, min_level/max_level i32, angle f32,
    desc u8[n,32], observations i32.  -> (nmatches, match_for_kp int32[n]: >=0 / -1 / -2)"""
    arrs = dict(valid=np.uint8, u=np.float32, v=np.float32, ur=np.float32, radius=np.float32, min_level=np.int32,
                max_level=np.int32, angle=np.float32, desc=np.uint8, observations=np.int32)
    keep = {k: np.ascontiguousarray(points[k], t) for k, t in arrs.items()}
    v = FramePointsView(len(keep["u"]), *[keep[k].ctypes.data for k in arrs])
    occ = np.ascontiguousarray(occupied, np.uint8)
    out = np.full(max(f[0].n, 1), -1, np.int32)
    nm = lib().orc_search_by_projection_frame(C.byref(f[0]), C.byref(v), int(check_ori), occ.ctypes.data, out.ctypes.data)
    return nm, out[:f[0].n]


def bow_view(desc, angle, feat_vec: dict, valid=None):
    """feat_vec: {node id: [feature indices]} (DBoW2::FeatureVector).  -> (BowView, keepalive)"""
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    angle = np.ascontiguousarray(angle, np.float32)
    ids = np.array(sorted(feat_vec), np.uint32)
    start = np.zeros(len(ids) + 1, np.int32)
    for i, k in enumerate(ids):
        start[i + 1] = start[i] + len(feat_vec[int(k)])
    idx = np.array([j for k in ids for j in feat_vec[int(k)]], np.uint32)
    keep = [desc, angle, ids, start, idx]
    v = BowView(len(desc), desc.ctypes.data, angle.ctypes.data, None, len(ids), ids.ctypes.data, start.ctypes.data, idx.ctypes.data)
    if valid is not None:
        va = np.ascontiguousarray(valid, np.uint8); keep.append(va); v.valid = va.ctypes.data
    return v, keep


def search_by_bow(s1, s2, th: int, nnratio: float, check_ori: bool):
    """-> (nmatches, m12 int32[n1], m21 int32[n2])"""
    m12 = np.empty(max(s1[0].n, 1), np.int32); m21 = np.empty(max(s2[0].n, 1), np.int32)
    nm = lib().orc_search_by_bow(C.byref(s1[0]), C.byref(s2[0]), th, nnratio, int(check_ori), m12.ctypes.data, m21.ctypes.data)
    return nm, m12[:s1[0].n], m21[:s2[0].n]


PROJ_POINT_FIELDS = dict(valid=np.uint8, u=np.float32, v=np.float32, ur=np.float32, radius=np.float32, level=np.int32,
                         angle=np.float32, desc=np.uint8)


def proj_points_view(points: dict):
    """points: dict of arrays valid u8, u / v / radius f32, level i32, desc u8[n,32]; ur and angle are
    optional (zeros).  -> (ProjPointsView, keepalive)"""
    n = len(points["u"])
    keep = {}
    for k, t in PROJ_POINT_FIELDS.items():
        keep[k] = np.ascontiguousarray(points[k], t) if k in points and points[k] is not None else np.zeros(max(n, 1), t)
    return ProjPointsView(n, *[keep[k].ctypes.data for k in PROJ_POINT_FIELDS]), keep


def search_by_projection_keyframe(f, points: dict, orb_dist: int, check_ori: bool, occupied: np.ndarray):
    """SearchByProjection(Frame&, KeyFrame*, ...) — -> (nmatches, match_for_kp int32[n]: >=0 / -1 / -2)"""
    v, keep = proj_points_view(points)
    occ = np.ascontiguousarray(occupied, np.uint8)
    out = np.full(max(f[0].n, 1), -1, np.int32)
    nm = lib().orc_sbp_frame_keyframe(C.byref(f[0]), C.byref(v), int(orb_dist), int(check_ori), occ.ctypes.data, out.ctypes.data)
    return nm, out[:f[0].n]


def search_by_projection_sim3(f, points: dict, matched: np.ndarray):
    """SearchByProjection(KeyFrame*, Scw, ...) — -> (nmatches, match_for_kp int32[n]: >=0 / -1)"""
    v, keep = proj_points_view(points)
    occ = np.ascontiguousarray(matched, np.uint8)
    out = np.full(max(f[0].n, 1), -1, np.int32)
    nm = lib().orc_sbp_keyframe_sim3(C.byref(f[0]), C.byref(v), occ.ctypes.data, out.ctypes.data)
    return nm, out[:f[0].n]


def fuse_search(f, points: dict, inv_level_sigma2=None):
    """search part of the two Fuse overloads -> (best_idx int32[n], best_dist int32[n])"""
    v, keep = proj_points_view(points)
    n = v.n
    bi = np.empty(max(n, 1), np.int32); bd = np.empty(max(n, 1), np.int32)
    sig = None if inv_level_sigma2 is None else np.ascontiguousarray(inv_level_sigma2, np.float32)
    lib().orc_fuse_search(C.byref(f[0]), C.byref(v), None if sig is None else sig.ctypes.data, bi.ctypes.data, bd.ctypes.data)
    return bi[:n], bd[:n]


def search_by_sim3(f1, f2, points1: dict, points2: dict):
    """-> (nFound, matches12 int32[n1])"""
    v1, k1 = proj_points_view(points1); v2, k2 = proj_points_view(points2)
    m12 = np.empty(max(v1.n, 1), np.int32)
    nf = lib().orc_search_by_sim3(C.byref(f1[0]), C.byref(f2[0]), C.byref(v1), C.byref(v2), m12.ctypes.data)
    return nf, m12[:v1.n]


def tri_side(kps, desc, has_mappoint, feat_vec: dict, uright=None):
    """One key frame of SearchForTriangulation -> (TriSide, keepalive)"""
    kps = np.ascontiguousarray(kps); desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    hm = np.ascontiguousarray(has_mappoint, np.uint8)
    ids = np.array(sorted(feat_vec), np.uint32)
    start = np.zeros(len(ids) + 1, np.int32)
    for i, k in enumerate(ids):
        start[i + 1] = start[i] + len(feat_vec[int(k)])
    idx = np.array([j for k in ids for j in feat_vec[int(k)]], np.uint32)
    keep = [kps, desc, hm, ids, start, idx]
    v = TriSide(len(kps), kps.ctypes.data, desc.ctypes.data, hm.ctypes.data, None, len(ids), ids.ctypes.data,
                start.ctypes.data, idx.ctypes.data)
    if uright is not None:
        ur = np.ascontiguousarray(uright, np.float32); keep.append(ur); v.uright = ur.ctypes.data
    return v, keep


def search_for_triangulation(s1, s2, F12, ex: float, ey: float, scale_factors2, level_sigma2_2, only_stereo=False,
                             check_ori=True):
    """-> (nmatches, matches12 int32[n1])"""
    F = np.ascontiguousarray(F12, np.float32).reshape(9)
    sf = np.ascontiguousarray(scale_factors2, np.float32); sg = np.ascontiguousarray(level_sigma2_2, np.float32)
    m12 = np.empty(max(s1[0].n, 1), np.int32)
    nm = lib().orc_search_for_triangulation(C.byref(s1[0]), C.byref(s2[0]), F.ctypes.data, float(ex), float(ey), sf.ctypes.data,
                                            sg.ctypes.data, int(only_stereo), int(check_ori), m12.ctypes.data)
    return nm, m12[:s1[0].n]


def distinctive_descriptors(desc: np.ndarray, start: np.ndarray) -> np.ndarray:
    """MapPoint::ComputeDistinctiveDescriptors for a batch: desc u8[total,32], start i32[npoints+1]."""
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32); start = np.ascontiguousarray(start, np.int32)
    best = np.empty(max(len(start) - 1, 1), np.int32)
    lib().orc_distinctive_descriptors(desc.ctypes.data, start.ctypes.data, len(start) - 1, best.ctypes.data)
    return best[:len(start) - 1]


def undistort_points(xy: np.ndarray, cam, dist) -> np.ndarray:
    """Frame::UndistortKeyPoints on (n, 2) float32 positions; cam = (fx, fy, cx, cy), dist = (k1, k2, p1, p2, k3)."""
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    cam = np.ascontiguousarray(cam, np.float32); dist = np.ascontiguousarray(dist, np.float32)
    assert cam.size == 4 and dist.size == 5
    out = np.empty_like(xy)
    lib().orc_undistort_points(xy.ctypes.data, len(xy), cam.ctypes.data, dist.ctypes.data, out.ctypes.data)
    return out


def image_bounds(cam, dist, cols: int, rows: int) -> np.ndarray:
    """Frame::ComputeImageBounds -> float32 (mnMinX, mnMaxX, mnMinY, mnMaxY)."""
    cam = np.ascontiguousarray(cam, np.float32); dist = np.ascontiguousarray(dist, np.float32)
    b = np.empty(4, np.float32)
    lib().orc_image_bounds(cam.ctypes.data, dist.ctypes.data, int(cols), int(rows), b.ctypes.data)
    return b


def assign_grid(f):
    """Frame::AssignFeaturesToGrid of a frame_view -> (cell_start int32[3073], items int32[k]); cell = ix*48+iy."""
    cs = np.zeros(64 * 48 + 1, np.int32); it = np.zeros(max(f[0].n, 1), np.int32)
    lib().orc_assign_grid(f[0], cs.ctypes.data, it.ctypes.data)
    return cs, it[:cs[-1]]


def pyramid_view(levels: list[np.ndarray]):
    n = len(levels)
    levels = [np.ascontiguousarray(l, np.uint8) for l in levels]
    ptrs = (C.c_void_p * n)(*[l.ctypes.data for l in levels])
    w = np.array([l.shape[1] for l in levels], np.int32); h = np.array([l.shape[0] for l in levels], np.int32)
    st = np.array([l.strides[0] for l in levels], np.int64)
    v = PyramidView(n, C.cast(ptrs, C.c_void_p), w.ctypes.data, h.ctypes.data, st.ctypes.data)
    return v, [levels, ptrs, w, h, st]


def compute_stereo_matches(fl, fr, pl, pr, inv_scale: np.ndarray, mbf: float, mb: float):
    n = fl[0].n
    ur = np.empty(max(n, 1), np.float32); dp = np.empty(max(n, 1), np.float32)
    inv = np.ascontiguousarray(inv_scale, np.float32)
    lib().orc_compute_stereo_matches(C.byref(fl[0]), C.byref(fr[0]), C.byref(pl[0]), C.byref(pr[0]), inv.ctypes.data, mbf, mb, ur.ctypes.data, dp.ctypes.data)
    return ur[:n], dp[:n]


def knn2(q: np.ndarray, t: np.ndarray, th=50, ratio=0.7):
    q = np.ascontiguousarray(q, np.uint8); t = np.ascontiguousarray(t, np.uint8)
    m = q.shape[0]
    idx = np.empty(m, np.int32); d1 = np.empty(m, np.int32); d2 = np.empty(m, np.int32)
    lib().orc_knn2(q.ctypes.data, m, t.ctypes.data, t.shape[0], th, ratio, idx.ctypes.data, d1.ctypes.data, d2.ctypes.data)
    return idx, d1, d2


# ----------------------------------------------------------------------------- the reference's own matcher TUs
# oracle/_ref/liborbref_matcher.so: UNMODIFIED src/ORBmatcher.cc, Frame.cc, KeyFrame.cc, MapPoint.cc, Map.cc (+ DBoW2)
# behind oracle/ref_matcher_wrap.cpp.  Same argument shapes as the restatement's wrappers above.
_mref = [False, None, "liborbref_matcher.so"]


def use_mref(name: str = "liborbref_matcher.so"):
    """Selects which build the ref_* wrappers below drive: the unmodified reference ("liborbref_matcher.so") or the
    drop-in proof ("liborbref_dropin.so": same TUs with ORBextractor.cc / ORBmatcher.cc replaced by integration/*.cc on
    liborbfe — needs a GPU)."""
    _mref[0], _mref[1], _mref[2] = False, None, name


def mref_lib():
    """None when the selected oracle/_ref library is not built (reference tree not mounted)."""
    if not _mref[0]:
        _mref[0] = True
        path = os.path.join(HERE, "_ref", _mref[2])
        if os.path.exists(path):
            L = C.CDLL(path)
            L.orm_search_for_initialization.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_int]
            L.orm_search_by_projection.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
            L.orm_search_by_projection_frame.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 8 + [C.c_float, C.c_int, C.c_float, C.c_int, C.c_void_p, C.c_void_p]
            L.orm_assign_grid.argtypes = [C.c_void_p] * 3
            L.orm_features_in_area.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_int]
            _mref[1] = L
    return _mref[1]


def ref_assign_grid(f):
    cs = np.empty(64 * 48 + 1, np.int32); items = np.empty(max(f[0].n, 1), np.int32)
    mref_lib().orm_assign_grid(C.byref(f[0]), cs.ctypes.data, items.ctypes.data)
    mref_lib().orm_reset()
    return cs, items[:f[0].n]


def ref_features_in_area(f, x, y, r, min_level=-1, max_level=-1):
    out = np.empty(max(f[0].n, 1), np.int32)
    n = mref_lib().orm_features_in_area(C.byref(f[0]), x, y, r, min_level, max_level, out.ctypes.data, len(out))
    mref_lib().orm_reset()
    return out[:n].copy()


def features_in_area(f, x, y, r, min_level=-1, max_level=-1):
    """the restatement's Frame / KeyFrame::GetFeaturesInArea on a view"""
    cs = np.empty(64 * 48 + 1, np.int32); items = np.empty(max(f[0].n, 1), np.int32)
    lib().orc_assign_grid(C.byref(f[0]), cs.ctypes.data, items.ctypes.data)
    out = np.empty(max(f[0].n, 1), np.int32)
    lib().orc_features_in_area.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_int]
    n = lib().orc_features_in_area(C.byref(f[0]), cs.ctypes.data, items.ctypes.data, x, y, r, min_level, max_level, out.ctypes.data, len(out))
    return out[:n].copy()


def ref_search_for_initialization(f1, f2, prev_matched: np.ndarray, window=100, nnratio=0.9, check_ori=True):
    n1 = f1[0].n
    m12 = np.empty(max(n1, 1), np.int32)
    nm = mref_lib().orm_search_for_initialization(C.byref(f1[0]), C.byref(f2[0]), prev_matched.ctypes.data, m12.ctypes.data, window, nnratio, int(check_ori))
    mref_lib().orm_reset()
    return nm, m12[:n1]


def ref_search_by_projection(f, mps: dict, th: float, nnratio: float, occupied: np.ndarray):
    arrs = dict(track_in_view=np.uint8, proj_x=np.float32, proj_y=np.float32, proj_xr=np.float32,
                level=np.int32, view_cos=np.float32, desc=np.uint8, observations=np.int32)
    keep = {k: np.ascontiguousarray(mps[k], t) for k, t in arrs.items()}
    v = MapPointsView(len(keep["level"]), *[keep[k].ctypes.data for k in
                      ("track_in_view", "proj_x", "proj_y", "proj_xr", "level", "view_cos", "desc", "observations")])
    occ = np.ascontiguousarray(occupied, np.uint8)
    out = np.full(max(f[0].n, 1), -1, np.int32)
    nm = mref_lib().orm_search_by_projection(C.byref(f[0]), C.byref(v), th, nnratio, occ.ctypes.data, out.ctypes.data)
    mref_lib().orm_reset()
    return nm, out[:f[0].n]


def ref_search_by_projection_frame(f, last: dict, th: float, mono: bool, bf: float, check_ori: bool, occupied: np.ndarray):
    """last: dict of arrays has_point / outlier u8, u / v f32, octave i32, angle f32, desc u8[n,32], observations i32
    (the last frame's keypoints with their map points at (u, v, 1)).  -> (nmatches, final index per keypoint or -1)"""
    arrs = dict(has_point=np.uint8, outlier=np.uint8, u=np.float32, v=np.float32, octave=np.int32, angle=np.float32, desc=np.uint8,
                observations=np.int32)
    keep = {k: np.ascontiguousarray(last[k], t) for k, t in arrs.items()}
    occ = np.ascontiguousarray(occupied, np.uint8)
    out = np.full(max(f[0].n, 1), -1, np.int32)
    nm = mref_lib().orm_search_by_projection_frame(C.byref(f[0]), len(keep["u"]), *[keep[k].ctypes.data for k in arrs], th, int(mono), bf,
                                                   int(check_ori), occ.ctypes.data, out.ctypes.data)
    mref_lib().orm_reset()
    return nm, out[:f[0].n]


def _mref_more():
    L = mref_lib()
    if getattr(L, "_more", False):
        return L
    P, I, F = C.c_void_p, C.c_int, C.c_float
    L.orm_search_by_bow_frame.argtypes = [P, P, F, I, P]
    L.orm_search_by_bow_keyframes.argtypes = [P, P, F, I, P]
    L.orm_sbp_frame_keyframe.argtypes = [P, P, P, P, P, F, I, I, P, P]
    L.orm_sbp_keyframe_sim3.argtypes = [P, P, P, P, I, P, P]
    L.orm_fuse.argtypes = [P, P, P, P, F, I, F, P, P, P, P]
    L.orm_search_by_sim3.argtypes = [P] * 9 + [F, P]
    L.orm_search_for_triangulation.argtypes = [P, P, P, F, F, P, I, F, F, I, I, P]
    L.orm_distinctive_descriptors.argtypes = [P, P, I, P]
    L.orm_compute_stereo_matches.argtypes = [P, P, P, P, I, I, F, F, P, P]
    L.orm_mono_frame.argtypes = [P, I, I, I, P, P, I, P, P, P, P, P, I]
    L.orm_stereo_frame.argtypes = [P, P, I, I, I, F, F, F, F, F, P, P, P, P, P, P, P, I]
    L._more = True
    return L


def _u8(a, n, default):
    return np.ascontiguousarray(a if a is not None else np.full(max(n, 1), default), np.uint8)


def ref_search_by_bow(s1, s2, nnratio: float, check_ori: bool, keyframes: bool):
    """keyframes False: SearchByBoW(KeyFrame*, Frame&, ..) -> (n, m21); True: SearchByBoW(KeyFrame*, KeyFrame*, ..) -> (n, m12)"""
    L = _mref_more()
    out = np.empty(max((s1 if keyframes else s2)[0].n, 1), np.int32)
    fn = L.orm_search_by_bow_keyframes if keyframes else L.orm_search_by_bow_frame
    n = fn(C.byref(s1[0]), C.byref(s2[0]), nnratio, int(check_ori), out.ctypes.data)
    L.orm_reset()
    return n, out[:(s1 if keyframes else s2)[0].n]


def ref_sbp_frame_keyframe(f, points: dict, has_point, bad, found, th, orb_dist, check_ori, occupied):
    L = _mref_more()
    v, keep = proj_points_view(points)
    n = v.n
    a = [_u8(has_point, n, 1), _u8(bad, n, 0), _u8(found, n, 0)]
    occ = np.ascontiguousarray(occupied, np.uint8)
    out = np.full(max(f[0].n, 1), -1, np.int32)
    nm = L.orm_sbp_frame_keyframe(C.byref(f[0]), C.byref(v), a[0].ctypes.data, a[1].ctypes.data, a[2].ctypes.data, th, orb_dist, int(check_ori),
                                  occ.ctypes.data, out.ctypes.data)
    L.orm_reset()
    return nm, out[:f[0].n]


def ref_sbp_keyframe_sim3(f, points: dict, bad, normal_ok, th: int, matched):
    L = _mref_more()
    v, keep = proj_points_view(points)
    a = [_u8(bad, v.n, 0), _u8(normal_ok, v.n, 1)]
    occ = np.ascontiguousarray(matched, np.uint8)
    out = np.full(max(f[0].n, 1), -1, np.int32)
    nm = L.orm_sbp_keyframe_sim3(C.byref(f[0]), C.byref(v), a[0].ctypes.data, a[1].ctypes.data, int(th), occ.ctypes.data, out.ctypes.data)
    L.orm_reset()
    return nm, out[:f[0].n]


def ref_fuse(f, points: dict, bad, normal_ok, th, with_pose: bool, bf=0.0, kp_has_point=None, kp_point_obs=None, point_obs=None):
    """-> (nFused, keypoint each point was fused with or -1)"""
    L = _mref_more()
    v, keep = proj_points_view(points)
    a = [_u8(bad, v.n, 0), _u8(normal_ok, v.n, 1), _u8(kp_has_point, f[0].n, 0)]
    ko = np.ascontiguousarray(kp_point_obs if kp_point_obs is not None else np.ones(max(f[0].n, 1)), np.int32)
    po = np.ascontiguousarray(point_obs if point_obs is not None else np.ones(max(v.n, 1)), np.int32)
    out = np.full(max(v.n, 1), -1, np.int32)
    nf = L.orm_fuse(C.byref(f[0]), C.byref(v), a[0].ctypes.data, a[1].ctypes.data, th, int(with_pose), bf, a[2].ctypes.data, ko.ctypes.data,
                    po.ctypes.data, out.ctypes.data)
    L.orm_reset()
    return nf, out[:v.n]


def ref_search_by_sim3(f1, f2, points1: dict, points2: dict, has1, has2, bad1, bad2, pre12, th):
    L = _mref_more()
    v1, k1 = proj_points_view(points1); v2, k2 = proj_points_view(points2)
    a = [_u8(has1, v1.n, 1), _u8(has2, v2.n, 1), _u8(bad1, v1.n, 0), _u8(bad2, v2.n, 0)]
    pre = np.ascontiguousarray(pre12 if pre12 is not None else np.full(max(v1.n, 1), -1), np.int32)
    out = np.empty(max(v1.n, 1), np.int32)
    n = L.orm_search_by_sim3(C.byref(f1[0]), C.byref(f2[0]), C.byref(v1), C.byref(v2), *[x.ctypes.data for x in a], pre.ctypes.data, th, out.ctypes.data)
    L.orm_reset()
    return n, out[:v1.n]


def ref_search_for_triangulation(s1, s2, F12, ex, ey, scale_factors, w, h, only_stereo=False, check_ori=True):
    L = _mref_more()
    Fm = np.ascontiguousarray(F12, np.float32).reshape(9)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    out = np.empty(max(s1[0].n, 1), np.int32)
    n = L.orm_search_for_triangulation(C.byref(s1[0]), C.byref(s2[0]), Fm.ctypes.data, ex, ey, sf.ctypes.data, len(sf), w, h, int(only_stereo),
                                       int(check_ori), out.ctypes.data)
    L.orm_reset()
    return n, out[:s1[0].n]


def ref_distinctive_descriptors(desc: np.ndarray, start: np.ndarray) -> np.ndarray:
    L = _mref_more()
    desc = np.ascontiguousarray(desc, np.uint8); start = np.ascontiguousarray(start, np.int32)
    best = np.empty(max(len(start) - 1, 1), np.int32)
    L.orm_distinctive_descriptors(desc.ctypes.data, start.ctypes.data, len(start) - 1, best.ctypes.data)
    L.orm_reset()
    return best[:len(start) - 1]


def ref_compute_stereo_matches(fl, fr, left: np.ndarray, right: np.ndarray, fx: float, bf: float):
    """Frame::ComputeStereoMatches of the reference on the given keypoint lists -> (mvuRight, mvDepth)"""
    L = _mref_more()
    h, w = left.shape
    left = np.ascontiguousarray(left); right = np.ascontiguousarray(right)
    ur = np.empty(max(fl[0].n, 1), np.float32); dp = np.empty(max(fl[0].n, 1), np.float32)
    L.orm_compute_stereo_matches(C.byref(fl[0]), C.byref(fr[0]), left.ctypes.data, right.ctypes.data, w, h, fx, bf, ur.ctypes.data, dp.ctypes.data)
    L.orm_reset()
    return ur[:fl[0].n], dp[:fl[0].n]


def ref_mono_frame(img: np.ndarray, nfeatures: int, cam, dist):
    """The reference's monocular Frame constructor -> (mvKeys, mvKeysUn, bounds[4], cell_start, cell_items)"""
    L = _mref_more()
    h, w = img.shape
    cap = 4 * nfeatures + 4096
    img = np.ascontiguousarray(img)
    k = np.empty(cap, KP_DTYPE); ku = np.empty(cap, KP_DTYPE); b = np.empty(4, np.float32)
    cs = np.empty(64 * 48 + 1, np.int32); items = np.empty(cap, np.int32)
    camv = np.ascontiguousarray(cam, np.float32); dv = np.ascontiguousarray(dist, np.float32)
    n = L.orm_mono_frame(img.ctypes.data, w, h, nfeatures, camv.ctypes.data, dv.ctypes.data, len(dv), k.ctypes.data, ku.ctypes.data, b.ctypes.data,
                         cs.ctypes.data, items.ctypes.data, cap)
    L.orm_reset()
    assert n >= 0
    return k[:n].copy(), ku[:n].copy(), b, cs, items[:cs[-1]].copy()


def ref_stereo_frame(left: np.ndarray, right: np.ndarray, nfeatures: int, cam, bf: float):
    """The reference's stereo Frame constructor (two extraction threads) -> (mvKeys, mDescriptors, mvuRight, mvDepth,
    mvKeysRight, mDescriptorsRight).  Meant for the drop-in build (see ref_matcher_wrap.cpp)."""
    L = _mref_more()
    h, w = left.shape
    cap = 4 * nfeatures + 4096
    k = np.empty(cap, KP_DTYPE); d = np.empty((cap, 32), np.uint8); ur = np.empty(cap, np.float32); dp = np.empty(cap, np.float32)
    kr = np.empty(cap, KP_DTYPE); dr = np.empty((cap, 32), np.uint8); nr = C.c_int(0)
    left = np.ascontiguousarray(left); right = np.ascontiguousarray(right)
    n = L.orm_stereo_frame(left.ctypes.data, right.ctypes.data, w, h, nfeatures, *[float(c) for c in cam], bf, k.ctypes.data, d.ctypes.data,
                           ur.ctypes.data, dp.ctypes.data, kr.ctypes.data, dr.ctypes.data, C.addressof(nr), cap)
    L.orm_reset()
    assert n >= 0
    return k[:n].copy(), d[:n].copy(), ur[:n].copy(), dp[:n].copy(), kr[:nr.value].copy(), dr[:nr.value].copy()


def dropin_compute_stereo_matches(fl, fr, left: np.ndarray, right: np.ndarray, fx: float, bf: float):
    """liborbref_dropin.so only: the adapter's ComputeStereoMatches(Frame&) on a reference Frame -> (mvuRight, mvDepth)"""
    L = _mref_more()
    L.orm_dropin_compute_stereo_matches.argtypes = [C.c_void_p] * 4 + [C.c_int, C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
    h, w = left.shape
    left = np.ascontiguousarray(left); right = np.ascontiguousarray(right)
    ur = np.empty(max(fl[0].n, 1), np.float32); dp = np.empty(max(fl[0].n, 1), np.float32)
    L.orm_dropin_compute_stereo_matches(C.byref(fl[0]), C.byref(fr[0]), left.ctypes.data, right.ctypes.data, w, h, fx, bf, ur.ctypes.data, dp.ctypes.data)
    L.orm_reset()
    return ur[:fl[0].n], dp[:fl[0].n]


def dropin_epilogue(img: np.ndarray, nfeatures: int, cam, dist):
    """liborbref_dropin.so only: the adapter's ComputeImageBounds + UndistortAndAssign on a Frame the reference's monocular
    constructor built -> (mvKeys, mvKeysUn, bounds[4], cell_start, cell_items)"""
    L = _mref_more()
    L.orm_dropin_epilogue.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int] + [C.c_void_p] * 5 + [C.c_int]
    h, w = img.shape
    cap = 4 * nfeatures + 4096
    img = np.ascontiguousarray(img)
    k = np.empty(cap, KP_DTYPE); ku = np.empty(cap, KP_DTYPE); b = np.empty(4, np.float32)
    cs = np.empty(64 * 48 + 1, np.int32); items = np.empty(cap, np.int32)
    camv = np.ascontiguousarray(cam, np.float32); dv = np.ascontiguousarray(dist, np.float32)
    n = L.orm_dropin_epilogue(img.ctypes.data, w, h, nfeatures, camv.ctypes.data, dv.ctypes.data, len(dv), k.ctypes.data, ku.ctypes.data, b.ctypes.data,
                              cs.ctypes.data, items.ctypes.data, cap)
    L.orm_reset()
    assert n >= 0
    return k[:n].copy(), ku[:n].copy(), b, cs, items[:cs[-1]].copy()
