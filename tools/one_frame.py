"""Extract one synthetic frame (debug helper for compute-sanitizer runs)."""
import numpy as np
from orb_slam2_annotate_b200 import synth
from orb_slam2_annotate_b200.extractor import ORBextractor
img = synth.frame(640, 480, 0)
ex = ORBextractor(1000, 1.2, 8, 20, 7)
k, d = ex(img)
print(len(k), d[:2])
