"""Pins the oracle's Frame-epilogue restatement (oracle/orc_frame.cpp): cv::undistortPoints as
Frame::UndistortKeyPoints calls it must reproduce the committed cv2 4.13 vectors bit for bit
(tests/golden/undistort_cv2.npz, made by tests/golden/make_golden.py), ComputeImageBounds
follows from the four corner points, and AssignFeaturesToGrid obeys PosInGrid's round()."""
import os

import numpy as np
import pytest

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "undistort_cv2.npz"))
CAMERAS = ("tum1", "tum2", "euroc", "wide")


@pytest.mark.parametrize("name", CAMERAS)
def test_undistort_points_matches_cv2(oracle, name):
    und = oracle.undistort_points(G[name + "_pts"], G[name + "_cam"], G[name + "_dist"])
    assert und.tobytes() == G[name + "_und"].tobytes()


@pytest.mark.parametrize("name", CAMERAS)
def test_image_bounds_from_cv2_corners(oracle, name):
    cols, rows = (int(v) for v in G[name + "_size"])
    c = G[name + "_und"][-4:]          # (0,0) (cols,0) (0,rows) (cols,rows), Frame.cc:485-488
    want = np.array([min(c[0, 0], c[2, 0]), max(c[1, 0], c[3, 0]), min(c[0, 1], c[1, 1]), max(c[2, 1], c[3, 1])], np.float32)
    got = oracle.image_bounds(G[name + "_cam"], G[name + "_dist"], cols, rows)
    assert got.tobytes() == want.tobytes()


def test_wide_lens_hits_the_negative_icdist_guard(oracle):
    # the corners of the strongly distorted camera come back unchanged (icdist < 0 branch)
    assert np.array_equal(G["wide_und"][-4:], G["wide_pts"][-4:])
    assert not np.array_equal(G["wide_und"][:3000], G["wide_pts"][:3000])


def test_no_k1_means_no_undistortion(oracle):
    pts = G["tum1_pts"]
    dist = G["tum1_dist"].copy(); dist[0] = 0.0         # only k1 is tested (Frame.cc:445)
    assert np.array_equal(oracle.undistort_points(pts, G["tum1_cam"], dist), pts)
    assert np.array_equal(oracle.image_bounds(G["tum1_cam"], dist, 640, 480), np.array([0, 640, 0, 480], np.float32))


def test_assign_grid_kat(oracle):
    b = (0.0, 640.0, 0.0, 480.0)                           # 10-px cells
    xy = np.array([[4.9, 4.9], [5.0, 5.0], [14.9, 0.0], [634.9, 474.9], [635.0, 10.0], [10.0, 475.0], [4.0, 4.0],
                   [-5.1, 3.0], [-4.9, 3.0]], np.float32)
    k = np.zeros(len(xy), oracle.KP_DTYPE); k["x"] = xy[:, 0]; k["y"] = xy[:, 1]
    f = oracle.frame_view(k, np.zeros((len(xy), 32), np.uint8), 640, 480, np.ones(8, np.float32), bounds=b)
    cs, items = oracle.assign_grid(f)
    cell = lambda ix, iy: list(items[cs[ix * 48 + iy]:cs[ix * 48 + iy + 1]])
    assert cell(0, 0) == [0, 6, 8]          # round(0.49) = 0; round(-0.49) = -0 -> cell 0
    assert cell(1, 1) == [1]                # round(0.5) = 1: half away from zero
    assert cell(1, 0) == [2]
    assert cell(63, 47) == [3]
    assert cs[-1] == 6                      # x -> 64, y -> 48 and x -> -1 fall outside (Frame.cc:426)
