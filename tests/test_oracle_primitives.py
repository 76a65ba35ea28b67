"""Pins the CPU oracle's OpenCV primitive restatements (oracle/orc_primitives.cpp) against
cv2 4.13.0: committed fixtures made by cv2 itself (tests/golden/prim_cv2.npz) and, where cv2 is
importable, live on fresh random inputs."""
import os

import numpy as np
import pytest

from orb_slam2_annotate_b200 import synth

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "prim_cv2.npz"))


@pytest.mark.parametrize("name", ["img", "noise"])
def test_primitives_against_cv2_fixtures(oracle, name):
    im = G[name]
    h, w = im.shape
    r = G[f"{name}_resize"]
    assert np.array_equal(oracle.resize_linear(im, r.shape[1], r.shape[0]), r)
    assert np.array_equal(oracle.copy_make_border(im, 19), G[f"{name}_border"])
    assert np.array_equal(oracle.gaussian_blur(im), G[f"{name}_blur"])
    for t in (20, 7):
        assert np.array_equal(oracle.fast(im, t, True), G[f"{name}_fast{t}"]), f"FAST t={t}"


def test_fast_atan2_fixture(oracle):
    yx, ref = G["atan2_yx"], G["atan2_deg"]
    got = np.array([oracle.fast_atan2(float(y), float(x)) for y, x in yx], np.float32)
    assert got.tobytes() == ref.tobytes()


def test_cv_round_half_even(oracle):
    L = oracle.lib()
    for v, e in [(0.5, 0), (1.5, 2), (2.5, 2), (-0.5, 0), (-1.5, -2), (3.4999, 3), (-2.5, -2), (1e6 + 0.5, 1000000)]:
        assert L.orc_cv_round_f(v) == e


def test_pattern_table_hash():
    import hashlib
    txt = open(os.path.join(os.path.dirname(__file__), "..", "include", "orbfe_pattern.inc")).read()
    body = txt[txt.index("*/") + 2:]
    vals = np.array([int(v) for v in body.replace("\n", "").split(",") if v.strip()], np.int32)
    assert vals.size == 1024 and vals.min() == -13 and vals.max() == 12
    # sha256 of the int32-LE table at reference src/ORBextractor.cc:155-413
    assert hashlib.sha256(vals.astype("<i4").tobytes()).hexdigest() == \
        "7e645581387b82784797e8adddb9b6f0c12611859fda09ca8a9bec96d767a05f"


def test_primitives_live_cv2(oracle):
    cv2 = pytest.importorskip("cv2")
    cv2.setNumThreads(1)
    rng = np.random.default_rng(7)
    for (w, h) in [(97, 61), (640, 480), (333, 77)]:
        for im in (rng.integers(0, 256, (h, w), dtype=np.uint8), synth.frame(w, h, 5)):
            dw, dh = int(round(w / 1.2)), int(round(h / 1.2))
            assert np.array_equal(oracle.resize_linear(im, dw, dh), cv2.resize(im, (dw, dh), interpolation=cv2.INTER_LINEAR))
            assert np.array_equal(oracle.gaussian_blur(im), cv2.GaussianBlur(im, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101))
            for t in (20, 7):
                det = cv2.FastFeatureDetector_create(t, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
                ref = np.array([[int(k.pt[0]), int(k.pt[1]), int(k.response)] for k in det.detect(im)], np.int32).reshape(-1, 3)
                assert np.array_equal(oracle.fast(im, t, True), ref)
