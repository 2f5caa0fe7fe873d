"""GPU parity of the input side of the pipeline: cvtColor(*2GRAY) and remap(INTER_LINEAR) fused
into the upload write the same level 0 as the oracle (itself pinned to cv2 4.13 by tests/golden/input_side.npz), and
extraction on top of them equals extraction of the oracle's converted / rectified image.  Bytes: bit-exact."""
import os

import numpy as np
import pytest

from my_orb_slam2_b200 import synth

pytestmark = pytest.mark.gpu
PARAMS = (600, 1.2, 8, 20, 7)


def _extractor(w, h, batch=1):
    from my_orb_slam2_b200 import orbfe
    return orbfe.Extractor(*PARAMS, w, h, batch)


@pytest.mark.parametrize("name,ch,rgb", [("rgb", 3, True), ("bgr", 3, False), ("rgba", 4, True), ("bgra", 4, False)])
def test_color_upload_golden(golden_dir, name, ch, rgb):
    g = np.load(os.path.join(golden_dir, "input_side.npz"))
    img = g["color_" + name]
    ex = _extractor(img.shape[1], img.shape[0])
    ex.upload_color(img[None], rgb)
    assert np.array_equal(ex.level(0), g["gray_" + name])


@pytest.mark.parametrize("ch,rgb,batch", [(3, True, 3), (4, False, 2)])
def test_color_upload_then_extract(oracle, ch, rgb, batch):
    w, h = 641, 363  # odd width: tail pixels of the 4-wide store path
    rng = np.random.default_rng(ch)
    gray = np.stack([synth.synth_image(w, h, seed=s) for s in range(batch)])
    # a colour image whose gray value is close to the synthetic scene, with coloured noise
    col = np.clip(gray[..., None].astype(np.int32) + rng.integers(-25, 26, (batch, h, w, ch)), 0, 255).astype(np.uint8)
    ex = _extractor(w, h, batch)
    ex.upload_color(col, rgb)
    ex.run(batch)
    kps = np.zeros((batch, ex.capacity), oracle.KEYPOINT_DTYPE)
    desc = np.zeros((batch, ex.capacity, 32), np.uint8)
    counts = np.zeros(batch, np.int32)
    ex.download(batch, kps, desc, counts)
    oex = oracle.Extractor(*PARAMS)
    for b in range(batch):
        want_gray = oracle.cvt_gray(col[b], rgb)
        assert np.array_equal(ex.level(0, b), want_gray)
        okp, odesc = oex.extract(want_gray)
        assert counts[b] == len(okp) > 300
        assert np.array_equal(kps[b][:counts[b]], okp) and np.array_equal(desc[b][:counts[b]], odesc)


@pytest.mark.parametrize("i", [0, 1])
def test_remap_upload_golden(golden_dir, i):
    g = np.load(os.path.join(golden_dir, "input_side.npz"))
    src, mx, my = g["remap_src_%d" % i], g["remap_mx_%d" % i], g["remap_my_%d" % i]
    ex = _extractor(mx.shape[1], mx.shape[0])
    ex.set_rectify_maps(mx, my, src.shape[1], src.shape[0])
    ex.upload_distorted(src[None])
    assert np.array_equal(ex.level(0), g["remap_dst_%d" % i])


def test_remap_upload_then_extract(oracle):
    """EuRoC-shaped: raw 752 x 480 frames rectified on the device, then extracted."""
    w, h = synth.SHAPES["euroc"]
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    r2 = ((xx - w / 2) ** 2 + (yy - h / 2) ** 2) / np.float32(w * w)
    mx = (xx + (xx - w / 2) * 0.12 * r2 + 1.37).astype(np.float32)  # barrel distortion + shift, leaves the image at the rim
    my = (yy + (yy - h / 2) * 0.12 * r2 - 0.61).astype(np.float32)
    raw = np.stack([synth.synth_image(w, h, seed=s) for s in (3, 4)])
    ex = _extractor(w, h, 2)
    ex.set_rectify_maps(mx, my, w, h)
    ex.upload_distorted(raw)
    ex.run(2)
    kps = np.zeros((2, ex.capacity), oracle.KEYPOINT_DTYPE)
    desc = np.zeros((2, ex.capacity, 32), np.uint8)
    counts = np.zeros(2, np.int32)
    ex.download(2, kps, desc, counts)
    oex = oracle.Extractor(*PARAMS)
    for b in range(2):
        rect = oracle.remap_linear(raw[b], mx, my)
        assert np.array_equal(ex.level(0, b), rect)
        okp, odesc = oex.extract(rect)
        assert counts[b] == len(okp) > 300
        assert np.array_equal(kps[b][:counts[b]], okp) and np.array_equal(desc[b][:counts[b]], odesc)


def test_distorted_without_maps_fails_loudly():
    from my_orb_slam2_b200 import orbfe
    ex = _extractor(64, 64)
    with pytest.raises(orbfe.OrbfeError):
        ex.upload_distorted(np.zeros((1, 64, 64), np.uint8))
