"""Single-frame latency through the C ABI (the shape Tracking uses: one 640x480 frame in, keypoints + descriptors out on
the host), and of SearchForInitialization on two such frames.  Prints one JSON line."""
import json
import time

import numpy as np

import orb_slam2_annotate_b200 as orbfe
from orb_slam2_annotate_b200 import synth

imgs, _ = synth.sequence(2, 640, 480, seed=4)
ex = orbfe.ORBextractor(1000, 1.2, 8, 20, 7)
mt = orbfe.ORBmatcher(0.9, True)
res = [ex(im) for im in imgs]
for _ in range(20):
    ex(imgs[0])
ts = []
for i in range(200):
    t0 = time.perf_counter(); ex(imgs[i & 1]); ts.append(time.perf_counter() - t0)
scale = ex.GetScaleFactors()
F = [orbfe.Frame(k, d, 640, 480, scale) for k, d in res]
prev = np.ascontiguousarray(np.stack([res[0][0]["x"], res[0][0]["y"]], 1), np.float32)
for _ in range(10):
    mt.SearchForInitialization(F[0], F[1], prev.copy(), 100)
tm = []
for _ in range(100):
    p = prev.copy()
    t0 = time.perf_counter(); mt.SearchForInitialization(F[0], F[1], p, 100); tm.append(time.perf_counter() - t0)
print(json.dumps({"extract_ms_median": float(np.median(ts)) * 1e3, "extract_ms_p90": float(np.percentile(ts, 90)) * 1e3,
                  "search_init_ms_median": float(np.median(tm)) * 1e3, "keypoints": int(len(res[0][0]))}))
