"""The C-ABI library loads on a CPU-only machine, exports every symbol include/orbfe.h
declares, and fails loudly (no CPU fallback) when asked to compute without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    h = open(os.path.join(ROOT, "include", "orbfe.h")).read()
    h = re.sub(r"/\*.*?\*/", "", h, flags=re.S)
    return sorted(set(re.findall(r"\b(orbfe_[a-z0-9_]+)\s*\(", h)))


def test_library_exports_every_declared_symbol(orbfe):
    from orb_slam2_annotate_b200 import _lib
    L = orbfe.lib()
    names = declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(L, n), f"liborbfe.so does not export {n}"
    assert set(_lib.SIGNATURES) == set(names), set(_lib.SIGNATURES) ^ set(names)
    assert L.orbfe_version().startswith(b"orbfe")


def test_keypoint_layout_is_cv_keypoint():
    from orb_slam2_annotate_b200 import KP_DTYPE
    assert KP_DTYPE.itemsize == 28
    assert [KP_DTYPE.fields[n][1] for n in ("x", "y", "size", "angle", "response", "octave", "class_id")] == [0, 4, 8, 12, 16, 20, 24]


def test_hamming256_host(orbfe):
    rng = np.random.default_rng(1)
    for _ in range(100):
        a = rng.integers(0, 256, 32, dtype=np.uint8); b = rng.integers(0, 256, 32, dtype=np.uint8)
        assert orbfe.ORBmatcher.DescriptorDistance(a, b) == int(np.unpackbits(a ^ b).sum())


def test_no_cpu_fallback(orbfe):
    L = orbfe.lib()
    if L.orbfe_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(orbfe.OrbfeError, match="no CUDA device"):
        orbfe.ORBextractor(1000)
    with pytest.raises(orbfe.OrbfeError, match="no CUDA device"):
        orbfe.ORBmatcher(0.7)
    h = C.c_void_p()
    assert L.orbfe_extractor_create(1000, 1.2, 8, 20, 7, 0, 0, 0, 0, C.byref(h)) == -4 and not h.value


def test_bad_arguments_rejected(orbfe):
    L = orbfe.lib()
    h = C.c_void_p()
    assert L.orbfe_extractor_create(0, 1.2, 8, 20, 7, 0, 0, 0, 0, C.byref(h)) == -1       # nfeatures
    assert L.orbfe_extractor_create(1000, 1.0, 8, 20, 7, 0, 0, 0, 0, C.byref(h)) == -1     # scale factor
    assert L.orbfe_extractor_create(1000, 1.2, 99, 20, 7, 0, 0, 0, 0, C.byref(h)) == -1    # levels
    assert L.orbfe_extractor_create(1000, 1.2, 8, 20, 7, 9000, 100, 1, 0, C.byref(h)) == -1
    assert b"" != L.orbfe_last_error()
