"""Oracle of the mono / RGB-D Frame constructors' tail (oracle/orb_frame.cc) against real cv2.undistortPoints output
(tests/golden/gen_golden_frame.py, python-cv2 4.13) and against plain numpy readings of src/Frame.cc:462-494, :688-712."""
import os

import numpy as np
import pytest


@pytest.fixture(scope="module")
def und(golden_dir):
    return np.load(os.path.join(golden_dir, "undistort.npz"))


def _kps(oracle, pts):
    k = np.zeros(len(pts), oracle.KEYPOINT_DTYPE)
    k["x"], k["y"] = pts[:, 0], pts[:, 1]
    k["size"], k["angle"], k["response"], k["octave"], k["class_id"] = 31.0, 12.5, 40.0, 1, -1
    return k


def test_undistort_keypoints_match_cv2(oracle, und):
    for name in und["names"]:
        cam, pts, want = und["cam_" + str(name)], und["pts_" + str(name)], und["und_" + str(name)]
        k = _kps(oracle, pts)
        u = oracle.undistort_keypoints(k, cam)
        if cam[4] == 0.0:   # mDistCoef.at<float>(0) == 0.0: mvKeysUn = mvKeys (src/Frame.cc:431-435)
            assert np.array_equal(u, k)
            continue
        assert np.array_equal(u["x"], want[:, 0]) and np.array_equal(u["y"], want[:, 1]), name
        for f in ("size", "angle", "response", "octave", "class_id"):
            assert np.array_equal(u[f], k[f])


def test_image_bounds_match_cv2_corners(oracle, und):
    for name in und["names"]:
        cam, c = und["cam_" + str(name)], und["corners_" + str(name)]
        b = oracle.image_bounds(cam, 640, 480)
        if cam[4] == 0.0:
            assert list(b) == [0.0, 640.0, 0.0, 480.0]
        else:  # src/Frame.cc:477-480
            want = [min(c[0, 0], c[2, 0]), max(c[1, 0], c[3, 0]), min(c[0, 1], c[1, 1]), max(c[2, 1], c[3, 1])]
            assert list(b) == [np.float32(v) for v in want], name


def test_stereo_from_rgbd_reading(oracle, und):
    rng = np.random.default_rng(5)
    pts = und["pts_tum1"][3000:5000]          # integer-valued keypoints inside the image
    k = _kps(oracle, pts)
    ku = oracle.undistort_keypoints(k, und["cam_tum1"])
    depth = rng.uniform(0.3, 8.0, (480, 640)).astype(np.float32)
    depth[rng.random((480, 640)) < 0.2] = 0.0    # holes of the depth camera
    depth[rng.random((480, 640)) < 0.02] = -1.0
    mbf = np.float32(40.0)
    ur, dz = oracle.stereo_from_rgbd(k, ku, depth, mbf)
    d = depth[k["y"].astype(np.int64), k["x"].astype(np.int64)]
    ok = d > 0
    assert np.array_equal(dz, np.where(ok, d, np.float32(-1)))
    with np.errstate(divide="ignore"):
        want = np.where(ok, ku["x"] - mbf / np.where(ok, d, np.float32(1)), np.float32(-1)).astype(np.float32)
    assert np.array_equal(ur, want)
    assert ok.sum() > 1000 and (~ok).sum() > 200
