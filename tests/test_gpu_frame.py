"""GPU parity of the tail of the mono / RGB-D Frame constructors (src/Frame.cc:118-228) through the C ABI:
orbfe_undistort_keypoints (Frame::UndistortKeyPoints, :429-460), orbfe_image_bounds (ComputeImageBounds, :462-494),
orbfe_stereo_from_rgbd (ComputeStereoFromRGBD, :688-712) against the oracle and the cv2-4.13 fixture: bit-exact."""
import os

import numpy as np
import pytest

from my_orb_slam2_b200 import synth

pytestmark = pytest.mark.gpu

TUM1 = (517.306408, 516.469215, 318.643040, 255.313989, 0.262383, -0.953104, -0.005358, 0.002628, 1.163314)


def _params(s):
    return (s["nfeatures"], s["scale_factor"], s["nlevels"], s["ini_th"], s["min_th"])


@pytest.mark.parametrize("cam", [TUM1, (535.4, 539.2, 320.1, 247.6, 0, 0, 0, 0, 0),
                                 (458.654, 457.296, 367.215, 248.375, -0.28340811, 0.07395907, 0.00019359, 1.76187114e-05)])
def test_mono_frame_tail_vs_oracle(oracle, cam):
    """TUM1-shaped frames (batch of 3): extract, then UndistortKeyPoints on the device == oracle on the same keypoints."""
    from my_orb_slam2_b200 import orbfe
    s = synth.SETTINGS["tum"]
    w, h = synth.SHAPES["tum"]
    imgs = np.stack([synth.synth_image(w, h, 70 + i) for i in range(3)])
    ex = orbfe.Extractor(*_params(s), w, h, max_batch=3)
    kps, descs, counts = ex.extract_batch(imgs)
    un = ex.undistort_keypoints(3, cam)
    for b in range(3):
        n = counts[b]
        want = oracle.undistort_keypoints(kps[b][:n], cam)
        assert np.array_equal(un[b][:n], want), "image %d" % b
        if cam[4] != 0:
            assert (un[b][:n]["x"] != kps[b][:n]["x"]).sum() > n // 2   # the distortion actually moved them
    assert np.array_equal(orbfe.image_bounds(cam, w, h), oracle.image_bounds(cam, w, h))


def test_undistort_matches_cv2_fixture_through_the_device(golden_dir):
    """The device evaluates orb_undistort_point with __dmul_rn / __dadd_rn / __ddiv_rn: compare it directly with the
    cv2 fixture by pushing fixture points through the kernel as the keypoints of an extractor run is not possible
    (keypoints come from images), so the comparison goes oracle == cv2 (CPU test) and device == oracle (above) — and here,
    on real extractor keypoints, device == the numpy model of cv2's loop evaluated in float64."""
    from my_orb_slam2_b200 import orbfe
    s = synth.SETTINGS["tum"]
    w, h = synth.SHAPES["tum"]
    ex = orbfe.Extractor(*_params(s), w, h, max_batch=1)
    kp, _ = ex.extract(synth.synth_image(w, h, 5))
    un = ex.undistort_keypoints(1, TUM1)[0][:len(kp)]
    fx, fy, cx, cy = (float(np.float32(v)) for v in TUM1[:4])
    k = [float(np.float32(v)) for v in TUM1[4:]]
    for i in range(0, len(kp), 7):
        x0 = (float(kp["x"][i]) - cx) * (1.0 / fx)
        y0 = (float(kp["y"][i]) - cy) * (1.0 / fy)
        x, y = x0, y0
        for _ in range(5):
            r2 = x * x + y * y
            ic = 1.0 / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2)
            dx = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x)
            dy = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y
            x, y = (x0 - dx) * ic, (y0 - dy) * ic
        assert un["x"][i] == np.float32(x * fx + cx) and un["y"][i] == np.float32(y * fy + cy)


def test_rgbd_frame_tail_vs_oracle(oracle):
    from my_orb_slam2_b200 import orbfe
    s = synth.SETTINGS["tum"]
    w, h = synth.SHAPES["tum"]
    rng = np.random.default_rng(9)
    imgs = np.stack([synth.synth_image(w, h, 80 + i) for i in range(2)])
    depth = rng.uniform(0.4, 9.0, (2, h, w)).astype(np.float32)
    depth[rng.random((2, h, w)) < 0.25] = 0.0
    depth[rng.random((2, h, w)) < 0.01] = -3.0
    mbf = np.float32(40.0)
    ex = orbfe.Extractor(*_params(s), w, h, max_batch=2)
    kps, descs, counts = ex.extract_batch(imgs)
    with pytest.raises(orbfe.OrbfeError):
        ex.stereo_from_rgbd(2, depth, mbf)          # needs mvKeysUn first
    un = ex.undistort_keypoints(2, TUM1)
    ur, dz = ex.stereo_from_rgbd(2, depth, mbf)
    for b in range(2):
        n = counts[b]
        our, odz = oracle.stereo_from_rgbd(kps[b][:n], un[b][:n], depth[b], mbf)
        assert np.array_equal(ur[b][:n], our) and np.array_equal(dz[b][:n], odz), "image %d" % b
        assert np.all(ur[b][n:] == -1) and np.all(dz[b][n:] == -1)
        assert (odz > 0).sum() > n // 2 and (odz < 0).sum() > n // 10
    ex.extract_batch(imgs)                           # a new extraction invalidates mvKeysUn
    with pytest.raises(orbfe.OrbfeError):
        ex.stereo_from_rgbd(2, depth, mbf)


def test_frontend_mono_and_rgbd_frames(oracle):
    """The Python mirror of the reference's mono / RGB-D Frame constructors (frontend.MonoFrame / RGBDFrame)."""
    from my_orb_slam2_b200 import frontend
    s = synth.SETTINGS["tum"]
    w, h = synth.SHAPES["tum"]
    img = synth.synth_image(w, h, 90)
    depth = np.random.default_rng(90).uniform(0.5, 6.0, (h, w)).astype(np.float32)
    ex = frontend.ORBextractor(*_params(s))
    f = frontend.RGBDFrame(img, depth, ex, TUM1[:4], TUM1[4:], 40.0)
    okp, odesc = oracle.Extractor(*_params(s)).extract(img)
    assert np.array_equal(f.mvKeys, okp) and np.array_equal(f.mDescriptors, odesc)
    oku = oracle.undistort_keypoints(okp, TUM1)
    assert np.array_equal(f.mvKeysUn, oku)
    our, odz = oracle.stereo_from_rgbd(okp, oku, depth, np.float32(40.0))
    assert np.array_equal(f.mvuRight, our) and np.array_equal(f.mvDepth, odz)
    assert np.array_equal(np.array([f.mnMinX, f.mnMaxX, f.mnMinY, f.mnMaxY], np.float32), oracle.image_bounds(TUM1, w, h))
