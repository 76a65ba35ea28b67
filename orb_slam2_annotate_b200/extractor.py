"""Python mirror of ``ORB_SLAM2::ORBextractor`` (reference include/ORBextractor.h:46-114) on top
of the C ABI.  Same constructor arguments, same call semantics, same getters; results are
numpy structured arrays laid out like ``cv::KeyPoint`` plus an ``N x 32`` uint8 descriptor
matrix.  All arithmetic runs in liborbfe's CUDA kernels; nothing here computes.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import KP_DTYPE, OrbfeError, check


class ORBextractor:
    """``ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)``.

    ``extractor(image, mask=None) -> (keypoints, descriptors)`` is ``operator()``
    (src/ORBextractor.cc:1119-1197): ``image`` is an HxW uint8 array; ``mask`` is accepted and
    ignored exactly as the reference ignores it.  An empty image returns ``(empty, None)``
    (the reference returns without touching its outputs, :1122-1123); zero keypoints also give
    ``descriptors is None`` (``_descriptors.release()``, :1145-1146).
    """

    def __init__(self, nfeatures: int = 1000, scaleFactor: float = 1.2, nlevels: int = 8,
                 iniThFAST: int = 20, minThFAST: int = 7, *, device: int = 0,
                 max_width: int = 0, max_height: int = 0, max_batch: int = 0):
        self._L = _lib.lib()
        self._h = C.c_void_p()
        self.nfeatures, self.nlevels = nfeatures, nlevels
        check(self._L.orbfe_extractor_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST,
                                             max_width, max_height, max_batch, device, C.byref(self._h)),
              "orbfe_extractor_create")
        n = nlevels
        self._scale, self._inv_scale, self._sigma2, self._inv_sigma2 = (np.empty(n, np.float32) for _ in range(4))
        self._per_level = np.empty(n, np.int32)
        self._umax = np.empty(16, np.int32)
        check(self._L.orbfe_get_scale_tables(self._h, self._scale.ctypes.data, self._inv_scale.ctypes.data,
                                             self._sigma2.ctypes.data, self._inv_sigma2.ctypes.data,
                                             self._per_level.ctypes.data, self._umax.ctypes.data))
        self._last_batch = 0
        self._bounds = None

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.orbfe_extractor_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # --- getters, include/ORBextractor.h:64-84
    def GetLevels(self) -> int:
        return self.nlevels

    def GetScaleFactor(self) -> float:
        return float(self._scale[1]) if self.nlevels > 1 else 1.0

    def GetScaleFactors(self) -> np.ndarray:
        return self._scale.copy()

    def GetInverseScaleFactors(self) -> np.ndarray:
        return self._inv_scale.copy()

    def GetScaleSigmaSquares(self) -> np.ndarray:
        return self._sigma2.copy()

    def GetInverseScaleSigmaSquares(self) -> np.ndarray:
        return self._inv_sigma2.copy()

    @property
    def features_per_level(self) -> np.ndarray:
        return self._per_level.copy()

    @property
    def umax(self) -> np.ndarray:
        return self._umax.copy()

    @property
    def stream(self) -> int:
        return int(self._L.orbfe_extractor_stream(self._h) or 0)

    @property
    def last_launches(self) -> int:
        return self._L.orbfe_extractor_last_launches(self._h)

    def max_keypoints(self) -> int:
        return self._L.orbfe_extractor_max_keypoints(self._h)

    # --- operator()
    def __call__(self, image: np.ndarray, mask=None):
        if image is None or image.size == 0:
            return np.empty(0, KP_DTYPE), None
        kps, desc, counts = self.extract_batch(image[None])
        n = int(counts[0])
        if n == 0:
            return np.empty(0, KP_DTYPE), None
        return kps[0, :n].copy(), desc[0, :n].copy()

    def extract_batch(self, images: np.ndarray, out=None):
        """images: [B,H,W] uint8 (rows contiguous).  Returns (kps [B,cap], desc [B,cap,32],
        counts [B]); only the first counts[b] slots of frame b are meaningful."""
        if images.dtype != np.uint8 or images.ndim != 3 or images.strides[2] != 1:
            raise ValueError("images must be a [B,H,W] uint8 array with contiguous rows")   # CV_8UC1 assert, :1126
        B, H, W = images.shape
        cap = self.max_keypoints()
        if out is None:
            out = (np.empty((B, cap), KP_DTYPE), np.empty((B, cap, 32), np.uint8), np.empty(B, np.int32))
        kps, desc, counts = out
        check(self._L.orbfe_extract_batch(self._h, images.ctypes.data, B, W, H, images.strides[1],
                                          images.strides[0], kps.ctypes.data, desc.ctypes.data,
                                          kps.shape[1], counts.ctypes.data), "orbfe_extract_batch")
        self._last_batch = B
        return kps, desc, counts

    def extract_batch_device(self, dev_ptr: int, B: int, W: int, H: int, stride: int | None = None,
                             frame_stride: int | None = None) -> None:
        """Enqueue extraction of frames already resident in HBM (asynchronous)."""
        stride = W if stride is None else stride
        frame_stride = stride * H if frame_stride is None else frame_stride
        check(self._L.orbfe_extract_batch_device(self._h, dev_ptr, B, W, H, stride, frame_stride),
              "orbfe_extract_batch_device")
        self._last_batch = B

    # --- Frame epilogue: UndistortKeyPoints + AssignFeaturesToGrid fused behind operator()
    def set_camera(self, camera, cols: int, rows: int) -> np.ndarray:
        """camera = (fx, fy, cx, cy, k1, k2, p1, p2[, k3]) or None to switch the epilogue off.
        Returns Frame's (mnMinX, mnMaxX, mnMinY, mnMaxY) for a cols x rows image
        (ComputeImageBounds, src/Frame.cc:481-512); afterwards every extract call also leaves
        mvKeysUn and mGrid of each frame on the device (``fetch_frames`` / ``device_frames``)."""
        bounds = np.zeros(4, np.float32)
        if camera is None:
            check(self._L.orbfe_extractor_set_camera(self._h, None, 0, 0, None))
            self._bounds = None
            return bounds
        cam = _lib.Camera(*(list(camera) + [0.0] * (9 - len(camera))))
        check(self._L.orbfe_extractor_set_camera(self._h, C.byref(cam), cols, rows, bounds.ctypes.data),
              "orbfe_extractor_set_camera")
        self._bounds = bounds
        return bounds.copy()

    def fetch_frames(self, B: int | None = None):
        """(kps_un [B,cap] KP_DTYPE, cell_end [B,3072] int32, items [B,cap] uint16) of the last batch:
        mGrid[ix][iy] of frame b = items[b, cell_end[b, c-1]:cell_end[b, c]] with c = ix*48 + iy."""
        B = self._last_batch if B is None else B
        cap = self.max_keypoints()
        kps_un = np.empty((B, cap), KP_DTYPE)
        cell_end = np.empty((B, _lib.GRID_CELLS), np.int32)
        items = np.empty((B, cap), np.uint16)
        check(self._L.orbfe_extractor_fetch_frames(self._h, B, kps_un.ctypes.data, cap, cell_end.ctypes.data,
                                                   items.ctypes.data), "orbfe_extractor_fetch_frames")
        return kps_un, cell_end, items

    def device_frames(self):
        """(kps_un_ptr, cell_end_ptr, items_ptr) device addresses of the last batch."""
        a, b, c = C.c_void_p(), C.c_void_p(), C.c_void_p()
        check(self._L.orbfe_extractor_device_frames(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def device_frame_view(self, frame: int, n: int, uright_ptr: int = 0):
        """``_lib.DeviceFrameView`` of frame ``frame`` of the last batch (n = its keypoint count) for the
        matcher's *_resident searches.  Keeps the scale-factor array alive on the returned object."""
        kp, ds, _, cap = self.device_results()
        ku, ce, it = self.device_frames()
        b = self._bounds
        v = _lib.DeviceFrameView(n, ku + frame * cap * KP_DTYPE.itemsize, ds + frame * cap * 32, uright_ptr or None,
                                 ce + frame * _lib.GRID_CELLS * 4, it + frame * cap * 2,
                                 float(b[0]), float(b[1]), float(b[2]), float(b[3]), self._scale.ctypes.data, self.nlevels)
        v._keep = self._scale
        return v

    def device_pyramid(self):
        """``_lib.DevicePyramidView`` of the last batch (mvImagePyramid of every frame, in device memory)."""
        v = _lib.DevicePyramidView()
        check(self._L.orbfe_extractor_device_pyramid(self._h, C.byref(v)), "orbfe_extractor_device_pyramid")
        return v

    STAGES = ("upload", "pyramid", "fast", "blur", "octree", "orient_desc", "frame")

    def set_profiling(self, on: bool = True) -> None:
        check(self._L.orbfe_extractor_set_profiling(self._h, int(on)))

    def stage_times_ms(self) -> dict:
        ms = np.zeros(len(self.STAGES), np.float32)
        check(self._L.orbfe_extractor_stage_times(self._h, ms.ctypes.data, len(ms)))
        return dict(zip(self.STAGES, ms.tolist()))

    def sync(self) -> None:
        check(self._L.orbfe_extractor_sync(self._h))

    def fetch(self, B: int | None = None, out=None):
        B = self._last_batch if B is None else B
        cap = self.max_keypoints()
        if out is None:
            out = (np.empty((B, cap), KP_DTYPE), np.empty((B, cap, 32), np.uint8), np.empty(B, np.int32))
        kps, desc, counts = out
        check(self._L.orbfe_extractor_fetch(self._h, B, kps.ctypes.data, desc.ctypes.data, kps.shape[1],
                                            counts.ctypes.data), "orbfe_extractor_fetch")
        return kps, desc, counts

    def device_results(self):
        """(kps_ptr, desc_ptr, counts_ptr, cap) device addresses of the last batch."""
        a, b, c, cap = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_int()
        check(self._L.orbfe_extractor_device_results(self._h, C.byref(a), C.byref(b), C.byref(c), C.byref(cap)))
        return a.value, b.value, c.value, cap.value

    # --- mvImagePyramid (include/ORBextractor.h:86), copied from the device on demand
    def pyramid_level(self, level: int, frame: int = 0, blurred: bool = False) -> np.ndarray:
        w, h = C.c_int(), C.c_int()
        check(self._L.orbfe_pyramid_level(self._h, frame, level, int(blurred), None, 0, C.byref(w), C.byref(h)))
        out = np.empty((h.value, w.value), np.uint8)
        check(self._L.orbfe_pyramid_level(self._h, frame, level, int(blurred), out.ctypes.data, out.strides[0],
                                          C.byref(w), C.byref(h)))
        return out

    @property
    def mvImagePyramid(self) -> list[np.ndarray]:
        return [self.pyramid_level(l) for l in range(self.nlevels)]

    def raw_keypoints(self, level: int, frame: int = 0) -> np.ndarray:
        """Per-level FAST corners before the quadtree, int32 [n,3] (x, y, score), reference order."""
        n = C.c_int()
        check(self._L.orbfe_level_raw_keypoints(self._h, frame, level, None, 0, C.byref(n)))
        out = np.empty((max(n.value, 1), 3), np.int32)
        check(self._L.orbfe_level_raw_keypoints(self._h, frame, level, out.ctypes.data, out.shape[0], C.byref(n)))
        return out[:n.value].copy()


def pinned_empty(shape, dtype=np.uint8) -> np.ndarray:
    """numpy array over page-locked host memory from orbfe_host_alloc (freed with the array)."""
    L = _lib.lib()
    dtype = np.dtype(dtype)
    nbytes = int(np.prod(shape)) * dtype.itemsize
    p = L.orbfe_host_alloc(max(nbytes, 1))
    if not p:
        raise OrbfeError("orbfe_host_alloc failed")
    buf = (C.c_uint8 * max(nbytes, 1)).from_address(p)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    _pinned_keepalive[arr.__array_interface__["data"][0]] = (p, buf)
    return arr


_pinned_keepalive: dict = {}


def pinned_free(arr: np.ndarray) -> None:
    key = arr.__array_interface__["data"][0]
    ent = _pinned_keepalive.pop(key, None)
    if ent:
        _lib.lib().orbfe_host_free(ent[0])
