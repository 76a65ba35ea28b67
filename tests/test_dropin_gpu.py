"""The drop-in proof on the GPU: oracle/_ref/liborbref_dropin.so is the reference's own Frame.cc, KeyFrame.cc, MapPoint.cc,
Map.cc and DBoW2 (unmodified) linked with integration/ORBextractor_orbfe.cc and integration/ORBmatcher_orbfe.cc — the
reference's class declarations implemented on liborbfe — in place of src/ORBextractor.cc and src/ORBmatcher.cc.  The
scenarios of tests/test_oracle_vs_ref_matcher.py are driven through it (reference objects in, reference call sites,
reference objects out) and must give what the unmodified library gave (tests/golden/ref_matcher.npz)."""
import os

import numpy as np
import pytest

import test_oracle_vs_ref_matcher as T

pytestmark = pytest.mark.gpu


@pytest.fixture()
def dropin(oracle):
    oracle.use_mref("liborbref_dropin.so")
    if oracle.mref_lib() is None:
        oracle.use_mref()
        pytest.skip("oracle/_ref/liborbref_dropin.so not built (reference tree not mounted at build time)")
    T._cache.pop(("epi", "tum1"), None); T._cache.pop(("epi", "tum2"), None); T._cache.pop(("epi", "plain"), None)
    yield oracle
    oracle.use_mref()
    for k in [k for k in T._cache if k[0] == "epi"]:
        T._cache.pop(k)


@pytest.mark.parametrize("scenario", [s for s in T.SCENARIOS if s is not T.scenario_frame_epilogue], ids=lambda s: s.__name__[9:])
def test_reference_call_sites_on_the_gpu(dropin, scenario):
    G = np.load(T.GOLDEN)
    got = scenario(dropin, True)
    for name, a in got.items():
        assert np.array_equal(a, G[name]), name


def test_frame_constructor_on_the_gpu(dropin):
    """Frame::Frame(imGray, ...) of the reference with the extraction served by liborbfe through Frame::ExtractORB: the
    keypoints equal the oracle extractor's, the epilogue (reference code on them) equals the restatement."""
    from orb_slam2_annotate_b200 import synth
    theirs = T.scenario_frame_epilogue(dropin, True)
    mine = T.scenario_frame_epilogue(dropin, False)
    for name, a in mine.items():
        assert np.array_equal(a, theirs[name]), name
    k = np.frombuffer(theirs["epi_tum1_keys"].tobytes(), dropin.KP_DTYPE)
    ko, _ = dropin.OracleExtractor(1000)(synth.frame(T.W, T.H, 50 + len("tum1")))
    assert len(k) == len(ko)
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(k[f], ko[f]), f
    assert np.array_equal(k["angle"], ko["angle"])


@pytest.mark.parametrize("w,h,nf,seed,disp,cam,bf", [(640, 480, 1000, 21, 12, (500.0, 500.0, 320.0, 240.0), 40.0),
                                                     (1241, 376, 2000, 2, 20, (718.856, 718.856, 607.1928, 185.2157), 386.1448)])
def test_stereo_frame_constructor_on_the_gpu(dropin, w, h, nf, seed, disp, cam, bf):
    """Frame::Frame(imLeft, imRight, ...) of the reference: its two extraction threads (src/Frame.cc:78-81) run two liborbfe
    handles concurrently through Frame::ExtractORB."""
    from orb_slam2_annotate_b200 import synth
    left, right = synth.stereo_pair(w, h, seed=seed, disparity=disp)
    for _ in range(3):                                    # the threads race for the device: the result must not depend on it
        k, d, ur, dp, kr, dr = dropin.ref_stereo_frame(left, right, nf, cam, bf)
        ol, orr = dropin.OracleExtractor(nf), dropin.OracleExtractor(nf)
        ko, do = ol(left); kro, dro = orr(right)
        for (a, b) in ((k, ko), (kr, kro)):
            assert len(a) == len(b)
            for f in ("x", "y", "size", "response", "octave", "class_id"):
                assert np.array_equal(a[f], b[f]), f
            assert np.array_equal(a["angle"], b["angle"])
        assert np.array_equal(d, do) and np.array_equal(dr, dro)
        # mvuRight / mvDepth of this constructor are not compared: it calls ComputeStereoMatches() (src/Frame.cc:90) before
        # `mb = mbf/fx` is assigned (:113), so the search range maxD = mbf/mb comes from an uninitialised member.  The routine
        # itself, entered with mb set, is compared in test_reference_call_sites_on_the_gpu[stereo].
        assert len(ur) == len(k) and len(dp) == len(k)


@pytest.mark.parametrize("w,h,nf,seed,disp,fx,bf", [(640, 480, 1000, 21, 12, 500.0, 40.0), (1241, 376, 2000, 2, 20, 718.856, 386.1448)])
def test_adapter_stereo_on_a_reference_frame(dropin, w, h, nf, seed, disp, fx, bf):
    """include/orbfe_adapter.hpp's ComputeStereoMatches(Frame&) — the replacement for Frame::ComputeStereoMatches — applied to the
    reference's own Frame object gives what the reference's member gives (fixture of the unmodified library)."""
    from orb_slam2_annotate_b200 import synth
    G = np.load(T.GOLDEN)
    left, right = synth.stereo_pair(w, h, seed=seed, disparity=disp)
    ol, orr = dropin.OracleExtractor(nf), dropin.OracleExtractor(nf)
    k, d = ol(left); kr, dr = orr(right)
    t = ol.tables()
    ur, dp = dropin.dropin_compute_stereo_matches(dropin.frame_view(k, d, w, h, t["scale"]), dropin.frame_view(kr, dr, w, h, t["scale"]), left, right, fx, bf)
    assert np.array_equal(ur.view(np.int32), G[f"stereo_{w}_uright"]) and np.array_equal(dp.view(np.int32), G[f"stereo_{w}_depth"])


def test_adapter_epilogue_on_a_reference_frame(dropin):
    """The adapter's ComputeImageBounds + UndistortAndAssign on a Frame built by the reference's
    monocular constructor reproduce what the constructor's own UndistortKeyPoints / ComputeImageBounds / AssignFeaturesToGrid left."""
    from orb_slam2_annotate_b200 import synth
    cams = {"tum1": ((517.306408, 516.469215, 318.643040, 255.313989), (0.262383, -0.953104, -0.005358, 0.002628, 1.163314)),
            "plain": ((500.0, 500.0, 320.0, 240.0), (0.0, 0.0, 0.0, 0.0))}
    for name, (cam, dist) in cams.items():
        img = synth.frame(T.W, T.H, 61)
        k, ku, b, cs, items = dropin.ref_mono_frame(img, 1000, cam, dist)
        k2, ku2, b2, cs2, items2 = dropin.dropin_epilogue(img, 1000, cam, dist)
        assert k.tobytes() == k2.tobytes()
        assert ku.tobytes() == ku2.tobytes(), name
        assert b.tobytes() == b2.tobytes() and np.array_equal(cs, cs2) and np.array_equal(items, items2), name
