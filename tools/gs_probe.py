"""Latency probe of the guided searches: direct C-ABI calls with prebuilt views (no numpy work per call)."""
import ctypes as C
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import orb_slam2_annotate_b200 as orbfe
from orb_slam2_annotate_b200 import _lib, synth

imgs, _ = synth.sequence(1, 640, 480, seed=1)
ex = orbfe.ORBextractor(1000)
ex.set_camera([520.0, 520.0, 320.0, 240.0, 0, 0, 0, 0, 0], 640, 480)
kps, desc, counts = ex.extract_batch(imgs)
n = int(counts[0]); k = kps[0, :n].copy(); d = desc[0, :n].copy()
scale = ex.GetScaleFactors()
rng = np.random.default_rng(4)
pts = dict(valid=np.ones(n, np.uint8), u=(k["x"] + rng.uniform(-4, 4, n)).astype(np.float32),
           v=(k["y"] + rng.uniform(-4, 4, n)).astype(np.float32), ur=np.zeros(n, np.float32),
           radius=(np.float32(15.0) * scale[k["octave"]]).astype(np.float32), min_level=(k["octave"] - 1).astype(np.int32),
           max_level=(k["octave"] + 1).astype(np.int32), angle=k["angle"].astype(np.float32),
           desc=d ^ rng.integers(0, 2, (n, 32), dtype=np.uint8), observations=rng.integers(0, 3, n).astype(np.int32))
mt = orbfe.ORBmatcher(0.8, True)
F = orbfe.Frame(k, d, 640, 480, scale)
fields = mt.FRAME_POINT_FIELDS
keep = {kk: np.ascontiguousarray(pts[kk], t) for kk, t in fields.items()}
v = _lib.FramePointsView(n, *[keep[kk].ctypes.data for kk in fields])
occ = np.zeros(n, np.uint8); out = np.full(n, -1, np.int32); nm = C.c_int()
L = _lib.lib()
fv = F.view(); dv = ex.device_frame_view(0, n)
for name, fn in (("host frame", lambda: L.orbfe_search_frame_points(mt._h, C.byref(fv), C.byref(v), 1, occ.ctypes.data, out.ctypes.data, C.byref(nm))),
                 ("resident frame", lambda: L.orbfe_search_frame_points_resident(mt._h, C.byref(dv), C.byref(v), 1, occ.ctypes.data, out.ctypes.data, C.byref(nm))),
                 ("python wrapper", lambda: mt.SearchByProjectionLast(dv, pts, occ))):
    for _ in range(5):
        fn()
    ts = []
    for _ in range(200):
        t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
    print(f"{name}: median {np.median(ts) * 1e6:.1f} us, min {np.min(ts) * 1e6:.1f} us, matches {nm.value}")
