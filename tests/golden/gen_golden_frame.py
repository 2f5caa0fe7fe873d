#!/usr/bin/env python3
"""Golden vectors for Frame::UndistortKeyPoints / ComputeImageBounds (src/Frame.cc:429-494): the real
cv2.undistortPoints(pts, K, D, None, K) of python-cv2 4.13 on the cameras of the reference's TUM settings files
(Examples/Monocular/TUM{1,2,3}.yaml: fx, fy, cx, cy, k1, k2, p1, p2, k3) and two more (four coefficients; strong
distortion).  Points: uniform over the image and a margin, integer-valued ones (level-0 keypoints), and level-scaled
ones (x * 1.2^l as the extractor produces them).

    python tests/golden/gen_golden_frame.py      # rewrites tests/golden/undistort.npz
"""
import os

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
CAMERAS = {
    "tum1": (517.306408, 516.469215, 318.643040, 255.313989, 0.262383, -0.953104, -0.005358, 0.002628, 1.163314),
    "tum2": (520.908620, 521.007327, 325.141442, 249.701764, 0.231222, -0.784899, -0.003257, -0.000105, 0.917205),
    "tum3": (535.4, 539.2, 320.1, 247.6, 0.0, 0.0, 0.0, 0.0, 0.0),            # no distortion: the copy branch
    "four": (458.654, 457.296, 367.215, 248.375, -0.28340811, 0.07395907, 0.00019359, 1.76187114e-05, 0.0),  # EuRoC cam0
    "strong": (300.0, 305.0, 310.0, 235.0, -0.45, 0.35, 0.002, -0.003, -0.12),
}


def main():
    assert cv2.__version__.startswith("4.13"), cv2.__version__
    rng = np.random.default_rng(2024)
    out = {"cv2_version": np.array(cv2.__version__), "names": np.array(list(CAMERAS))}
    for name, cam in CAMERAS.items():
        K = np.array([[cam[0], 0, cam[2]], [0, cam[1], cam[3]], [0, 0, 1]], np.float32)
        D = np.array(cam[4:], np.float32)
        a = np.stack([rng.uniform(-30, 670, 3000), rng.uniform(-30, 510, 3000)], 1)
        b = np.round(np.stack([rng.uniform(19, 620, 2000), rng.uniform(19, 460, 2000)], 1))
        lv = rng.integers(1, 8, 2000)
        c = np.round(np.stack([rng.uniform(19, 500, 2000), rng.uniform(19, 380, 2000)], 1)) * (np.float32(1.2) ** lv)[:, None]
        pts = np.concatenate([a, b, c]).astype(np.float32)
        und = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, D, None, K).reshape(-1, 2)
        corners = np.array([[0, 0], [640, 0], [0, 480], [640, 480]], np.float32)
        cu = cv2.undistortPoints(corners.reshape(-1, 1, 2), K, D, None, K).reshape(-1, 2)
        out["cam_" + name] = np.array(cam, np.float32)
        out["pts_" + name] = pts
        out["und_" + name] = und
        out["corners_" + name] = cu
    path = os.path.join(HERE, "undistort.npz")
    np.savez_compressed(path, **out)
    print("undistort ->", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
