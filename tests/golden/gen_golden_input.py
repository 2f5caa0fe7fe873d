#!/usr/bin/env python3
"""Golden vectors for the input side of the pipeline, made with python cv2 (OpenCV 4.13.0):
cv2.cvtColor(..., COLOR_{RGB,BGR,RGBA,BGRA}2GRAY) as Tracking::GrabImage* calls it (src/Tracking.cc:189-214) and
cv2.remap(src, mapx, mapy, INTER_LINEAR) with CV_32F maps as Examples/Stereo/stereo_euroc.cc:136-137 calls it.

    python tests/golden/gen_golden_input.py        # rewrites tests/golden/input_side.npz
"""
import os

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def rectify_maps(w, h, seed):
    """CV_32F maps of a mildly distorted EuRoC-like camera (cv2.initUndistortRectifyMap itself)."""
    rng = np.random.default_rng(seed)
    K = np.array([[458.654 * w / 752, 0, 367.215 * w / 752], [0, 457.296 * h / 480, 248.375 * h / 480], [0, 0, 1]])
    D = np.array([-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05]) * rng.uniform(0.8, 1.2)
    ang = 0.01 * rng.normal(size=3)
    R, _ = cv2.Rodrigues(ang)
    P = K.copy()
    P[0, 2] += 3.0
    return cv2.initUndistortRectifyMap(K, D, R, P, (w, h), cv2.CV_32F)


def main():
    rng = np.random.default_rng(7)
    out = {}
    for name, ch, code in (("rgb", 3, cv2.COLOR_RGB2GRAY), ("bgr", 3, cv2.COLOR_BGR2GRAY),
                           ("rgba", 4, cv2.COLOR_RGBA2GRAY), ("bgra", 4, cv2.COLOR_BGRA2GRAY)):
        img = rng.integers(0, 256, (45, 67, ch), dtype=np.uint8)
        img[0, :8] = [[0] * ch, [255] * ch, [255, 0, 0, 9][:ch], [0, 255, 0, 9][:ch], [0, 0, 255, 9][:ch],
                      [1, 1, 1, 0][:ch], [254, 255, 254, 0][:ch], [128, 127, 129, 0][:ch]]
        out["color_" + name] = img
        out["gray_" + name] = cv2.cvtColor(img, code)
    for i, (w, h) in enumerate([(94, 60), (188, 120)]):
        src = rng.integers(0, 256, (h, w), dtype=np.uint8)
        mx, my = rectify_maps(w, h, i)
        if i == 1:  # exercise the border: push part of the map outside the source, add exact half-way cases
            mx = mx.copy(); my = my.copy()
            mx[:10] -= 40.0
            my[:, -12:] += 30.0
            mx[20, :32] = np.arange(32, dtype=np.float32) / 64.0 + 5.0   # multiples of 1/64: cvRound ties
            my[20, :32] = 7.0 + np.arange(32, dtype=np.float32) / 64.0
        out["remap_src_%d" % i] = src
        out["remap_mx_%d" % i] = mx
        out["remap_my_%d" % i] = my
        out["remap_dst_%d" % i] = cv2.remap(src, mx, my, cv2.INTER_LINEAR)
    np.savez_compressed(os.path.join(HERE, "input_side.npz"), **out)
    print("wrote input_side.npz with", sorted(out))


if __name__ == "__main__":
    main()
