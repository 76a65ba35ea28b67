"""Writes tests/golden/ref_matcher.npz: the outputs of the UNMODIFIED reference matcher-side translation units
(oracle/_ref/liborbref_matcher.so — src/ORBmatcher.cc, Frame.cc, KeyFrame.cc, MapPoint.cc, Map.cc, DBoW2) on the seeded
scenarios of tests/test_oracle_vs_ref_matcher.py.  Needs the reference tree (run `make -C oracle` first); the fixture lets
the restatement be checked where the reference is not mounted.
usage: python tests/golden/make_ref_matcher_golden.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import oracle_py as oracle                    # noqa: E402
import test_oracle_vs_ref_matcher as T                     # noqa: E402

if __name__ == "__main__":
    oracle.build()
    assert oracle.mref_lib() is not None, "oracle/_ref/liborbref_matcher.so is not built"
    out = {}
    for sc in T.SCENARIOS:
        got = sc(oracle, True)
        assert not set(got) & set(out)
        out.update(got)
    np.savez_compressed(T.GOLDEN, **out)
    print("wrote", T.GOLDEN, len(out), "arrays,", os.path.getsize(T.GOLDEN), "bytes")
