"""ctypes loader for liborbfe.so (the C ABI declared in include/orbfe.h).

The library is built in-tree by ``orb_slam2_annotate_b200/csrc/Makefile`` (see
``__graft_entry__.build``).  There is no CPU implementation behind this package: if the
shared library is missing, or no CUDA device is present, every compute entry point raises.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liborbfe.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28


class OrbfeError(RuntimeError):
    pass


class FrameView(C.Structure):
    _fields_ = [("n", C.c_int32), ("kps", C.c_void_p), ("desc", C.c_void_p), ("uright", C.c_void_p),
                ("min_x", C.c_float), ("max_x", C.c_float), ("min_y", C.c_float), ("max_y", C.c_float),
                ("scale_factors", C.c_void_p), ("nlevels", C.c_int32), ("keyframe_bounds", C.c_int32),
                ("grid_inv_w", C.c_float), ("grid_inv_h", C.c_float)]


class MapPointsView(C.Structure):
    _fields_ = [("n", C.c_int32), ("track_in_view", C.c_void_p), ("proj_x", C.c_void_p),
                ("proj_y", C.c_void_p), ("proj_xr", C.c_void_p), ("level", C.c_void_p),
                ("view_cos", C.c_void_p), ("desc", C.c_void_p), ("observations", C.c_void_p)]


class FramePointsView(C.Structure):
    _fields_ = [("n", C.c_int32), ("valid", C.c_void_p), ("u", C.c_void_p), ("v", C.c_void_p),
                ("ur", C.c_void_p), ("radius", C.c_void_p), ("min_level", C.c_void_p),
                ("max_level", C.c_void_p), ("angle", C.c_void_p), ("desc", C.c_void_p),
                ("observations", C.c_void_p)]


class Camera(C.Structure):
    """mK entries + mDistCoef (k1 k2 p1 p2 k3) as Frame holds them."""
    _fields_ = [(k, C.c_float) for k in ("fx", "fy", "cx", "cy", "k1", "k2", "p1", "p2", "k3")]


class DeviceFrameView(C.Structure):
    _fields_ = [("n", C.c_int32), ("d_kps", C.c_void_p), ("d_desc", C.c_void_p), ("d_uright", C.c_void_p),
                ("d_cell_end", C.c_void_p), ("d_items", C.c_void_p),
                ("min_x", C.c_float), ("max_x", C.c_float), ("min_y", C.c_float), ("max_y", C.c_float),
                ("scale_factors", C.c_void_p), ("nlevels", C.c_int32), ("keyframe_bounds", C.c_int32),
                ("grid_inv_w", C.c_float), ("grid_inv_h", C.c_float)]


GRID_COLS, GRID_ROWS = 64, 48
GRID_CELLS = GRID_COLS * GRID_ROWS


class BowView(C.Structure):
    _fields_ = [("n", C.c_int32), ("desc", C.c_void_p), ("angle", C.c_void_p), ("valid", C.c_void_p),
                ("nnodes", C.c_int32), ("node_id", C.c_void_p), ("node_start", C.c_void_p), ("indices", C.c_void_p)]


class WindowPointsView(C.Structure):
    _fields_ = [("n", C.c_int32), ("valid", C.c_void_p), ("u", C.c_void_p), ("v", C.c_void_p),
                ("ur", C.c_void_p), ("radius", C.c_void_p), ("level", C.c_void_p), ("desc", C.c_void_p)]


class TriSideView(C.Structure):
    _fields_ = [("n", C.c_int32), ("kps", C.c_void_p), ("desc", C.c_void_p), ("has_mappoint", C.c_void_p),
                ("uright", C.c_void_p), ("nnodes", C.c_int32), ("node_id", C.c_void_p), ("node_start", C.c_void_p),
                ("indices", C.c_void_p)]


class DevicePyramidView(C.Structure):
    _fields_ = [("nlevels", C.c_int32), ("d_data", C.c_void_p), ("width", C.c_void_p), ("height", C.c_void_p),
                ("pitch", C.c_void_p), ("frame_stride", C.c_void_p)]


class PyramidView(C.Structure):
    _fields_ = [("nlevels", C.c_int32), ("data", C.c_void_p), ("width", C.c_void_p),
                ("height", C.c_void_p), ("stride", C.c_void_p)]


def build(force: bool = False) -> None:
    """Compile liborbfe.so for sm_100a with nvcc (cross-compiles without a GPU)."""
    if force:
        subprocess.run(["make", "-C", os.path.join(HERE, "csrc"), "clean"], check=True, stdout=subprocess.DEVNULL)
    subprocess.run(["make", "-C", os.path.join(HERE, "csrc"), "-j8", "all"], check=True, stdout=subprocess.DEVNULL)


_lib = None

_vp, _i, _f, _ip = C.c_void_p, C.c_int, C.c_float, C.POINTER(C.c_int)
_ss = C.c_ssize_t

# name -> (restype, argtypes); every symbol include/orbfe.h declares
SIGNATURES = {
    "orbfe_last_error": (C.c_char_p, []),
    "orbfe_version": (C.c_char_p, []),
    "orbfe_device_count": (_i, []),
    "orbfe_host_alloc": (_vp, [C.c_size_t]),
    "orbfe_host_free": (None, [_vp]),
    "orbfe_extractor_create": (_i, [_i, _f, _i, _i, _i, _i, _i, _i, _i, C.POINTER(_vp)]),
    "orbfe_extractor_destroy": (None, [_vp]),
    "orbfe_get_scale_tables": (_i, [_vp] + [_vp] * 6),
    "orbfe_extractor_levels": (_i, [_vp]),
    "orbfe_extractor_max_keypoints": (_i, [_vp]),
    "orbfe_extract": (_i, [_vp, _vp, _i, _i, _ss, _vp, _vp, _i, _ip]),
    "orbfe_extract_batch": (_i, [_vp, _vp, _i, _i, _i, _ss, _ss, _vp, _vp, _i, _vp]),
    "orbfe_extract_batch_device": (_i, [_vp, _vp, _i, _i, _i, _ss, _ss]),
    "orbfe_extractor_sync": (_i, [_vp]),
    "orbfe_extractor_device_results": (_i, [_vp, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp), _ip]),
    "orbfe_extractor_fetch": (_i, [_vp, _i, _vp, _vp, _i, _vp]),
    "orbfe_extractor_stream": (_vp, [_vp]),
    "orbfe_extractor_last_launches": (_i, [_vp]),
    "orbfe_extractor_set_profiling": (_i, [_vp, _i]),
    "orbfe_extractor_stage_times": (_i, [_vp, _vp, _i]),
    "orbfe_pyramid_level": (_i, [_vp, _i, _i, _i, _vp, _ss, _ip, _ip]),
    "orbfe_pyramid_host": (_i, [_vp, _i, _vp, _vp, _vp, _vp]),
    "orbfe_level_raw_keypoints": (_i, [_vp, _i, _i, _vp, _i, _ip]),
    "orbfe_hamming256": (_i, [_vp, _vp]),
    "orbfe_matcher_create": (_i, [_i, C.POINTER(_vp)]),
    "orbfe_matcher_destroy": (None, [_vp]),
    "orbfe_matcher_stream": (_vp, [_vp]),
    "orbfe_matcher_sync": (_i, [_vp]),
    "orbfe_knn2": (_i, [_vp, _vp, _i, _vp, _i, _i, _f, _vp, _vp, _vp]),
    "orbfe_knn2_device": (_i, [_vp, _vp, _i, _vp, _i, _i, _f, _vp, _vp, _vp]),
    "orbfe_matcher_fetch": (_i, [_vp, _vp, _vp, C.c_size_t]),
    "orbfe_matcher_set_knn_engine": (_i, [_vp, _i]),
    "orbfe_knn2_device_noalloc": (_i, [_i, _vp, _vp, _i, _vp, _i, _i, _f, _vp, _vp, _vp]),
    "orbfe_comm_unique_id": (_i, [_vp]),
    "orbfe_comm_create": (_i, [_vp, _i, _i, _i, C.POINTER(_vp)]),
    "orbfe_comm_wrap": (_i, [_vp, _i, _i, _i, C.POINTER(_vp)]),
    "orbfe_comm_destroy": (None, [_vp]),
    "orbfe_comm_world": (_i, [_vp]),
    "orbfe_comm_rank": (_i, [_vp]),
    "orbfe_knn2_shard_rows": (_i, [_i, _i]),
    "orbfe_knn2_sharded": (_i, [_vp, _vp, _vp, _i, _vp, _i, _i, _f, _vp]),
    "orbfe_search_init": (_i, [_vp, C.POINTER(FrameView), C.POINTER(FrameView), _vp, _vp, _i, _f, _i, _ip]),
    "orbfe_search_init_consecutive_device": (_i, [_vp, _vp, _vp, _vp, _i, _i, _f, _f, _f, _f, _i, _f, _i, _vp, _vp]),
    "orbfe_search_local_points": (_i, [_vp, C.POINTER(FrameView), C.POINTER(MapPointsView), _f, _f, _vp, _vp, _ip]),
    "orbfe_search_frame_points": (_i, [_vp, C.POINTER(FrameView), C.POINTER(FramePointsView), _i, _vp, _vp, _ip]),
    "orbfe_extractor_set_camera": (_i, [_vp, C.POINTER(Camera), _i, _i, _vp]),
    "orbfe_extractor_device_frames": (_i, [_vp, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp)]),
    "orbfe_extractor_fetch_frames": (_i, [_vp, _i, _vp, _i, _vp, _vp]),
    "orbfe_image_bounds": (_i, [_vp, C.POINTER(Camera), _i, _i, _vp]),
    "orbfe_frame_assign": (_i, [_vp, C.POINTER(Camera), _vp, _vp, _i, _vp, _vp, _vp, _ip]),
    "orbfe_search_local_points_resident": (_i, [_vp, C.POINTER(DeviceFrameView), C.POINTER(MapPointsView), _f, _f, _vp, _vp, _ip]),
    "orbfe_search_frame_points_resident": (_i, [_vp, C.POINTER(DeviceFrameView), C.POINTER(FramePointsView), _i, _vp, _vp, _ip]),
    "orbfe_search_projected_points": (_i, [_vp, C.POINTER(FrameView), C.POINTER(FramePointsView), _i, _i, _vp, _vp, _ip]),
    "orbfe_search_projected_points_resident": (_i, [_vp, C.POINTER(DeviceFrameView), C.POINTER(FramePointsView), _i, _i, _vp, _vp, _ip]),
    "orbfe_search_window_best": (_i, [_vp, C.POINTER(FrameView), C.POINTER(WindowPointsView), _vp, _i, _vp, _vp]),
    "orbfe_search_window_best_resident": (_i, [_vp, C.POINTER(DeviceFrameView), C.POINTER(WindowPointsView), _vp, _i, _vp, _vp]),
    "orbfe_search_by_sim3": (_i, [_vp, C.POINTER(FrameView), C.POINTER(FrameView), C.POINTER(WindowPointsView),
                                  C.POINTER(WindowPointsView), _i, _vp, _ip]),
    "orbfe_search_triangulation": (_i, [_vp, C.POINTER(TriSideView), C.POINTER(TriSideView), _vp, _f, _f, _vp, _vp, _i, _i, _i,
                                        _vp, _ip]),
    "orbfe_search_bow": (_i, [_vp, C.POINTER(BowView), C.POINTER(BowView), _i, _f, _i, _vp, _vp, _ip]),
    "orbfe_search_bow_batch": (_i, [_vp, _i, _vp, _vp, _i, _f, _i, _vp, _vp, _vp]),
    "orbfe_distinctive_descriptors": (_i, [_vp, _vp, _vp, _i, _vp]),
    "orbfe_extractor_device_pyramid": (_i, [_vp, C.POINTER(DevicePyramidView)]),
    "orbfe_stereo_match_batch_device": (_i, [_vp, C.POINTER(DevicePyramidView), _vp, _vp, _vp, _i, _i, _vp, _vp, _i, _f, _f, _vp, _vp]),
    "orbfe_stereo_match": (_i, [_vp, C.POINTER(FrameView), C.POINTER(FrameView), C.POINTER(PyramidView),
                                C.POINTER(PyramidView), _vp, _f, _f, _vp, _vp]),
}


def lib() -> C.CDLL:
    """Load liborbfe.so; raises OrbfeError when it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise OrbfeError(f"{LIB_PATH} is missing: run __graft_entry__.build() "
                             "(make -C orb_slam2_annotate_b200/csrc); there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)   # AttributeError here means the .so is stale
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        raise OrbfeError(f"{what or 'orbfe call'} failed ({rc}): {lib().orbfe_last_error().decode()}")
