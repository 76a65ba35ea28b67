"""Times the reference's monocular Frame constructor (src/Frame.cc:173-243) and SearchForInitialization on reference
objects, once in the unmodified build (oracle/_ref/liborbref_matcher.so) and once in the drop-in build
(liborbref_dropin.so: integration/*.cc on liborbfe).  Prints one JSON line.  usage: PYTHONPATH=. python tools/dropin_probe.py"""
import json
import time

import numpy as np

from oracle import oracle_py as O
from orb_slam2_annotate_b200 import synth

imgs, _ = synth.sequence(2, 640, 480, seed=4)
cam, dist = (517.306408, 516.469215, 318.643040, 255.313989), (0.262383, -0.953104, -0.005358, 0.002628, 1.163314)
out = {}
for name, lib in (("reference", "liborbref_matcher.so"), ("dropin", "liborbref_dropin.so")):
    O.use_mref(lib)
    if O.mref_lib() is None:
        continue
    O.ref_mono_frame(imgs[0], 1000, cam, dist)
    n = 5 if name == "reference" else 200
    t0 = time.perf_counter()
    for i in range(n):
        k, ku, b, cs, items = O.ref_mono_frame(imgs[i & 1], 1000, cam, dist)
    out[name + "_frame_ctor_ms"] = (time.perf_counter() - t0) / n * 1e3
    oc = O.OracleExtractor(1000)
    fr = [oc(im) for im in imgs]
    t = oc.tables()
    f1, f2 = (O.frame_view(k, d, 640, 480, t["scale"]) for k, d in fr)
    prev = np.ascontiguousarray(np.stack([fr[0][0]["x"], fr[0][0]["y"]], 1), np.float32)
    O.ref_search_for_initialization(f1, f2, prev.copy(), 100, 0.9, True)
    m = 20 if name == "reference" else 200
    import ctypes
    O.mref_lib().orm_last_call_ms.restype = ctypes.c_double
    t0 = time.perf_counter()
    inner = []
    for _ in range(m):
        O.ref_search_for_initialization(f1, f2, prev.copy(), 100, 0.9, True)
        inner.append(O.mref_lib().orm_last_call_ms())
    out[name + "_search_init_ms"] = (time.perf_counter() - t0) / m * 1e3             # with the harness building two Frames
    out[name + "_search_init_call_ms"] = float(np.median(inner))                     # the SearchForInitialization call alone
O.use_mref()
print(json.dumps(out))
