"""The C++ host side (include/orbfe_adapter.hpp: ORB_SLAM2::ORBextractor / ORBmatcher with the
reference's signatures) compiles against a cv:: type shim and, on the GPU box, produces the same
keypoints, descriptors and initialisation matches as the Python host side and the oracle."""
import os
import subprocess

import numpy as np
import pytest

from orb_slam2_annotate_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "cpp", "adapter_smoke")


def _build():
    subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "liborb_oracle.so"], check=True, stdout=subprocess.DEVNULL)
    subprocess.run(["g++", "-std=c++11", "-O2", "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "oracle", "cvshim"),
                    "-I" + os.path.join(ROOT, "oracle"), "-o", EXE, os.path.join(ROOT, "tests", "cpp", "adapter_smoke.cpp"),
                    "-L" + os.path.join(ROOT, "orb_slam2_annotate_b200"), "-lorbfe", "-L" + os.path.join(ROOT, "oracle"), "-lorb_oracle",
                    "-Wl,-rpath," + os.path.join(ROOT, "orb_slam2_annotate_b200"), "-Wl,-rpath," + os.path.join(ROOT, "oracle")],
                   check=True)


def test_adapter_compiles_and_links(orbfe):
    _build()
    assert os.path.exists(EXE)


@pytest.mark.gpu
def test_adapter_end_to_end(orbfe, oracle, tmp_path):
    _build()
    imgs, _ = synth.sequence(2, 640, 480, seed=4)
    inp, out = str(tmp_path / "in.bin"), str(tmp_path / "out.bin")
    with open(inp, "wb") as f:
        f.write(np.array([640, 480], np.int32).tobytes()); f.write(imgs.tobytes())
    r = subprocess.run([EXE, inp, out], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    print(r.stdout.strip())
    raw = open(out, "rb").read()
    off = 0
    frames = []
    for i in range(2):
        n = int(np.frombuffer(raw, np.int32, 1, off)[0]); off += 4
        k = np.frombuffer(raw, orbfe.KP_DTYPE, n, off); off += 28 * n
        d = np.frombuffer(raw, np.uint8, 32 * n, off).reshape(n, 32); off += 32 * n
        frames.append((k, d))
    nm, n1 = (int(v) for v in np.frombuffer(raw, np.int32, 2, off)); off += 8
    m12 = np.frombuffer(raw, np.int32, n1, off); off += 4 * n1
    t = oracle.OracleExtractor(2000).tables()
    ref = [oracle.OracleExtractor(2000)(im) for im in imgs]
    for (k, d), (ko, do) in zip(frames, ref):
        assert len(k) == len(ko)
        for f in ("x", "y", "size", "response", "octave", "class_id"):
            assert np.array_equal(k[f], ko[f])
        assert np.abs(k["angle"] - ko["angle"]).max() <= 1e-3 and (d != do).any(axis=1).sum() <= 2
    prev = np.ascontiguousarray(np.stack([ref[0][0]["x"], ref[0][0]["y"]], 1), np.float32)
    no, mo = oracle.search_for_initialization(oracle.frame_view(ref[0][0], ref[0][1], 640, 480, t["scale"]),
                                              oracle.frame_view(ref[1][0], ref[1][1], 640, 480, t["scale"]), prev, 100, 0.9, True)
    assert nm == no and np.array_equal(m12, mo) and nm > 100
    # Frame epilogue through the adapter (TUM1 camera): bounds, mvKeysUn and mGrid of frame 0
    cam = np.array([517.306408, 516.469215, 318.643040, 255.313989], np.float32)
    dist = np.array([0.262383, -0.953104, -0.005358, 0.002628, 1.163314], np.float32)
    bnd = np.frombuffer(raw, np.float32, 4, off); off += 16
    assert bnd.tobytes() == oracle.image_bounds(cam, dist, 640, 480).tobytes()
    k0 = frames[0][0]
    un = np.frombuffer(raw, np.float32, 2 * len(k0), off).reshape(-1, 2); off += 8 * len(k0)
    want = oracle.undistort_points(np.stack([k0["x"], k0["y"]], 1), cam, dist)
    assert un.tobytes() == want.tobytes()
    ku = k0.copy(); ku["x"], ku["y"] = want[:, 0], want[:, 1]
    cs, items = oracle.assign_grid(oracle.frame_view(ku, frames[0][1], 640, 480, t["scale"], bounds=bnd))
    for c in range(64 * 48):
        cnt = int(np.frombuffer(raw, np.int32, 1, off)[0]); off += 4
        got = np.frombuffer(raw, np.int32, cnt, off); off += 4 * cnt
        assert np.array_equal(got, items[cs[c]:cs[c + 1]])
    # both SearchByBoW overloads through the adapter: counts against the oracle on the same stand-in feature vectors
    nb1, cnt1, len1, nb2, cnt2, len2 = (int(v) for v in np.frombuffer(raw, np.int32, 6, off)); off += 24
    assert off == len(raw)
    sides = []
    for k, d in frames:
        fv = {}
        for i in range(len(k)):
            fv.setdefault(int((d[i, 0] & 15) | (k["octave"][i] << 4)), []).append(i)
        idx = np.arange(len(k))
        valid = ((idx % 3 != 0) & ~((idx % 7 == 0) & (idx % 3 != 0))).astype(np.uint8)
        sides.append((d, k["angle"], fv, valid))
    (da, aa, fva, va), (db, ab, fvb, vb) = sides
    n1, _, o21 = oracle.search_by_bow(oracle.bow_view(da, aa, fva, va), oracle.bow_view(db, ab, fvb), 50, 0.75, True)
    n2, o12, _ = oracle.search_by_bow(oracle.bow_view(da, aa, fva, va), oracle.bow_view(db, ab, fvb, vb), 49, 0.75, True)
    assert (nb1, cnt1, len1) == (n1, int((o21 >= 0).sum()), len(db)) and n1 > 30
    assert (nb2, cnt2, len2) == (n2, int((o12 >= 0).sum()), len(da)) and n2 > 10
