"""Where the single-frame call spends its time: pageable vs pinned host buffers vs device-resident frames."""
import json, os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import orb_slam2_annotate_b200 as orbfe
from orb_slam2_annotate_b200 import synth

imgs, _ = synth.sequence(2, 640, 480, seed=4)
ex = orbfe.ORBextractor(1000, 1.2, 8, 20, 7)
cap = ex.max_keypoints()

def med(fn, n=300, warm=30):
    for _ in range(warm): fn()
    ts = []
    for _ in range(n):
        t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
    return float(np.median(ts)) * 1e3

out = {}
out["pageable_ms"] = med(lambda: ex(imgs[0]))
pin = orbfe.pinned_empty((1, 480, 640), np.uint8); pin[0] = imgs[0]
pout = (orbfe.pinned_empty((1, cap), orbfe.KP_DTYPE), orbfe.pinned_empty((1, cap, 32), np.uint8), orbfe.pinned_empty((1,), np.int32))
out["pinned_ms"] = med(lambda: ex.extract_batch(pin, out=pout))
dev = torch.from_numpy(imgs[:1].copy()).cuda()
def resident():
    ex.extract_batch_device(dev.data_ptr(), 1, 640, 480); ex.sync()
out["device_resident_ms"] = med(resident)
def resident_fetch():
    ex.extract_batch_device(dev.data_ptr(), 1, 640, 480); ex.fetch(1, out=pout)
out["device_resident_plus_fetch_ms"] = med(resident_fetch)
h = torch.from_numpy(pin)
def h2d_only():
    dev.copy_(h, non_blocking=True); torch.cuda.synchronize()
out["h2d_307KB_pinned_ms"] = med(h2d_only)
print(json.dumps(out))
