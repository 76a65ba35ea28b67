"""Pins the oracle's extractor restatement (oracle/orc_extractor.cpp): against committed
outputs of the UNMODIFIED reference TU (tests/golden/ref_extract.npz, produced by
oracle/_ref/liborbref.so) and, where oracle/_ref is built, live on further frames."""
import hashlib
import os

import numpy as np
import pytest

from orb_slam2_annotate_b200 import synth

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_extract.npz"))
CASES = sorted(k[:-4] for k in G.files if k.endswith("_cfg"))


@pytest.mark.parametrize("case", CASES)
def test_oracle_equals_reference_fixture(oracle, case):
    w, h, nf, seed = (int(v) for v in G[case + "_cfg"])
    img = synth.frame(w, h, seed)
    assert hashlib.sha256(img.tobytes()).hexdigest() == str(G[case + "_img_sha"]), "synthetic frame generator drifted"
    oc = oracle.OracleExtractor(nf)
    k, d = oc(img)
    assert k.tobytes() == G[case + "_kps"].tobytes()
    assert d.tobytes() == G[case + "_desc"].tobytes()
    for l in range(8):
        assert hashlib.sha256(oc.level(l).tobytes()).hexdigest() == str(G[case + "_pyr_sha"][l])


def test_oracle_tables(oracle):
    t = oracle.OracleExtractor(1000).tables()
    assert t["features_per_level"].tolist() == [217, 181, 151, 126, 105, 87, 73, 60]
    assert oracle.OracleExtractor(2000).tables()["features_per_level"].tolist() == [434, 362, 302, 251, 209, 175, 145, 122]
    assert t["umax"].tolist() == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]   # :471-485
    assert np.allclose(t["scale"], 1.2 ** np.arange(8), rtol=1e-6)


def test_oracle_equals_reference_live(oracle):
    if oracle.ref_lib("") is None:
        pytest.skip("oracle/_ref not built (reference tree not mounted)")
    for (w, h, nf, seed) in [(640, 480, 1000, 11), (752, 480, 1200, 12), (1241, 376, 2000, 13)]:
        img = synth.frame(w, h, seed)
        ka, da = oracle.OracleExtractor(nf)(img)
        r = oracle.RefExtractor(nf)
        kr, dr = r(img)
        assert ka.tobytes() == kr.tobytes() and da.tobytes() == dr.tobytes()
        assert np.array_equal(r.tables()["scale"], oracle.OracleExtractor(nf).tables()["scale"])
    # degenerate frames
    for img in (np.full((480, 640), 9, np.uint8), synth.noise_frame(320, 240, 2)):
        ka, da = oracle.OracleExtractor(500)(img)
        kr, dr = oracle.RefExtractor(500)(img)
        assert ka.tobytes() == kr.tobytes() and da.tobytes() == dr.tobytes()


def test_quadtree_allocator_dependence_reported(oracle):
    """The reference breaks quadtree size ties by heap address (ORBextractor.cc:728): the same
    TU on glibc malloc may select different keypoints.  Reported, not graded."""
    if oracle.ref_lib("_malloc") is None:
        pytest.skip("oracle/_ref not built")
    img = synth.frame(640, 480, 0)
    kb, _ = oracle.RefExtractor(1000)(img)
    km, _ = oracle.RefExtractor(1000, variant="_malloc")(img)
    sb = {(float(a), float(b), int(o)) for a, b, o in zip(kb["x"], kb["y"], kb["octave"])}
    sm = {(float(a), float(b), int(o)) for a, b, o in zip(km["x"], km["y"], km["octave"])}
    print(f"bump-allocator contract vs glibc malloc: {len(sb & sm)} of {len(sb)} keypoints in common")
    assert len(sb & sm) > 0.5 * len(sb)
