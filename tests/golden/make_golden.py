"""Generates the committed golden fixtures (run in the BUILD container, where cv2 4.13.0 and
/root/reference are available; the GPU box has neither).

  prim_*.npz      inputs + outputs of the OpenCV primitives the reference calls, produced by
                  cv2 4.13.0 itself (resize INTER_LINEAR, copyMakeBorder REFLECT_101, FAST 9/16
                  with NMS at thresholds 20 and 7, GaussianBlur 7x7 sigma 2, fastAtan2)
  undistort_cv2.npz  cv2.undistortPoints as Frame::UndistortKeyPoints calls it, for the reference's
                  shipped TUM1 / TUM2 / EuRoC cameras and a wide-angle one
  ref_extract.npz keypoints + descriptors of the UNMODIFIED reference TU (oracle/_ref/liborbref.so,
                  src/ORBextractor.cc compiled against oracle/cvshim, bump allocator) on seeded
                  synthetic frames, plus per-level pyramid checksums
Usage: python tests/golden/make_golden.py
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from orb_slam2_annotate_b200 import synth  # noqa: E402


def primitives():
    import cv2
    cv2.setNumThreads(1)
    rng = np.random.default_rng(123)
    img = synth.frame(160, 120, 3)
    noise = rng.integers(0, 256, (64, 80), dtype=np.uint8)
    out = {"cv2_version": np.array(cv2.__version__), "img": img, "noise": noise}
    for name, im in (("img", img), ("noise", noise)):
        h, w = im.shape
        dw, dh = int(round(w / 1.2)), int(round(h / 1.2))
        out[f"{name}_resize"] = cv2.resize(im, (dw, dh), interpolation=cv2.INTER_LINEAR)
        out[f"{name}_border"] = cv2.copyMakeBorder(im, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
        out[f"{name}_blur"] = cv2.GaussianBlur(im, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
        for t in (20, 7):
            det = cv2.FastFeatureDetector_create(t, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
            kps = det.detect(im)
            out[f"{name}_fast{t}"] = np.array([[int(k.pt[0]), int(k.pt[1]), int(k.response)] for k in kps], np.int32).reshape(-1, 3)
    yx = rng.integers(-1_400_000, 1_400_000, (4096, 2)).astype(np.float32)
    yx[:16] = [[0, 1], [1, 0], [0, -1], [-1, 0], [1, 1], [-1, 1], [1, -1], [-1, -1],
               [0, 0], [3, 4], [-3, 4], [5, -12], [1e-3, 1], [1, 1e-3], [-7, -7], [2, 1]]
    out["atan2_yx"] = yx
    out["atan2_deg"] = np.array([cv2.fastAtan2(float(y), float(x)) for y, x in yx], np.float32)
    np.savez_compressed(os.path.join(HERE, "prim_cv2.npz"), **out)
    print("prim_cv2.npz written with cv2", cv2.__version__)


def reference_extract():
    from oracle import oracle_py as O
    O.build()
    if O.ref_lib("") is None:
        raise SystemExit("oracle/_ref/liborbref.so missing: /root/reference not mounted?")
    out = {}
    cases = [("A", 640, 480, 1000, 0), ("A", 640, 480, 1000, 1), ("A2", 640, 480, 2000, 0), ("B", 1241, 376, 2000, 0),
             ("S", 321, 243, 500, 4)]
    for name, w, h, nf, seed in cases:
        img = synth.frame(w, h, seed)
        ex = O.RefExtractor(nf)
        k, d = ex(img)
        key = f"{name}_{seed}"
        out[key + "_cfg"] = np.array([w, h, nf, seed], np.int32)
        out[key + "_kps"] = k
        out[key + "_desc"] = d
        out[key + "_pyr_sha"] = np.array([hashlib.sha256(ex.level(l).tobytes()).hexdigest() for l in range(8)])
        out[key + "_img_sha"] = np.array(hashlib.sha256(img.tobytes()).hexdigest())
    np.savez_compressed(os.path.join(HERE, "ref_extract.npz"), **out)
    print("ref_extract.npz written:", [k for k in out if k.endswith("_cfg")])


# the reference's shipped camera files (Examples/Monocular/TUM1.yaml, TUM2.yaml, EuRoC.yaml)
# plus a strongly distorted wide-angle lens; (fx, fy, cx, cy), (k1, k2, p1, p2, k3), cols, rows
CAMERAS = {
    "tum1": ((517.306408, 516.469215, 318.643040, 255.313989), (0.262383, -0.953104, -0.005358, 0.002628, 1.163314), 640, 480),
    "tum2": ((520.908620, 521.007327, 325.141442, 249.701764), (0.231222, -0.784899, -0.003257, -0.000105, 0.917205), 640, 480),
    "euroc": ((458.654, 457.296, 367.215, 248.375), (-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05, 0.0), 752, 480),
    "wide": ((300.0, 300.0, 320.0, 240.0), (-0.45, 0.3, 0.01, -0.02, -0.1), 640, 480),
}


def undistort():
    """cv::undistortPoints exactly as Frame::UndistortKeyPoints / ComputeImageBounds call it
    (Frame.cc:463, :493): CV_32FC2 points, float K and distCoef, R empty, P = K."""
    import cv2
    rng = np.random.default_rng(7)
    out = {"cv2_version": np.array(cv2.__version__)}
    for name, (cam, dist, cols, rows) in CAMERAS.items():
        fx, fy, cx, cy = cam
        K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
        D = np.array(dist, np.float32).reshape(-1, 1)
        # keypoint-like positions (integer pixel * level scale) and arbitrary sub-pixel ones
        scale = (1.2 ** rng.integers(0, 8, 3000)).astype(np.float32)
        kp = (rng.integers(16, [cols - 16, rows - 16], (3000, 2)).astype(np.float32) / scale[:, None]).round() * scale[:, None]
        pts = np.concatenate([kp.astype(np.float32), rng.uniform([-40, -40], [cols + 40, rows + 40], (1000, 2)).astype(np.float32),
                              np.array([[0, 0], [cols, 0], [0, rows], [cols, rows]], np.float32)])
        und = cv2.undistortPoints(pts.reshape(-1, 1, 2).copy(), K, D, None, K).reshape(-1, 2)
        out[name + "_cam"] = np.array(cam, np.float32)
        out[name + "_dist"] = np.array(dist, np.float32)
        out[name + "_size"] = np.array([cols, rows], np.int32)
        out[name + "_pts"] = pts
        out[name + "_und"] = und.astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "undistort_cv2.npz"), **out)
    print("undistort_cv2.npz written with cv2", cv2.__version__)


if __name__ == "__main__":
    primitives()
    undistort()
    reference_extract()
