"""GPU parity of the extractor path (C ABI -> sm_100a kernels) against the CPU oracle and the cv2 golden fixtures.
Bit-exact bar for coordinates, octaves, responses, descriptors and pyramid bytes; angles are compared bit-exact too
(the tolerance north_star allows is 1e-3 degrees — see ANGLE_TOL — and any descriptor difference caused by an angle
flip would be counted by test_angle_tolerance_report)."""
import os

import numpy as np
import pytest

from my_orb_slam2_b200 import synth

pytestmark = pytest.mark.gpu
ANGLE_TOL = 1e-3  # degrees (BASELINE.json north_star)


@pytest.fixture(scope="module")
def fe():
    from my_orb_slam2_b200 import frontend, orbfe
    assert orbfe.device_count() >= 1, "no CUDA device: the product has no CPU fallback"
    return frontend


def _assert_same(kp, desc, okp, odesc):
    assert len(kp) == len(okp)
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(kp[f], okp[f]), f
    assert np.max(np.abs(kp["angle"] - okp["angle"]), initial=0) <= ANGLE_TOL
    assert np.array_equal(kp["angle"], okp["angle"])
    assert np.array_equal(desc, odesc)


@pytest.mark.parametrize("name", ["small_a", "small_b", "small_noise"])
def test_golden_fixtures(fe, golden_dir, name):
    g = np.load(os.path.join(golden_dir, "extract_%s.npz" % name))
    p = g["params"]
    ex = fe.ORBextractor(int(p[0]), float(p[1]), int(p[2]), int(p[3]), int(p[4]))
    kp, desc = ex(g["image"])
    gk = g["keypoints"]
    assert len(kp) == len(gk)
    assert np.array_equal(kp["x"], gk[:, 0]) and np.array_equal(kp["y"], gk[:, 1])
    assert np.array_equal(kp["size"], gk[:, 2]) and np.array_equal(kp["angle"], gk[:, 3])
    assert np.array_equal(kp["response"], gk[:, 4]) and np.array_equal(kp["octave"], gk[:, 5].astype(np.int32))
    assert np.array_equal(desc, g["descriptors"])
    assert np.array_equal(ex.mvImagePyramid[int(p[2]) - 1], g["level_last"])


@pytest.mark.parametrize("shape", ["tum", "kitti", "euroc"])
def test_full_size_golden_fixtures(fe, golden_dir, shape):
    """CUDA path against the committed cv2-4.13 chain at full size (tests/golden/gen_golden.py): keypoints, descriptors,
    every pyramid level by sha256."""
    import hashlib
    from my_orb_slam2_b200 import synth
    g = np.load(os.path.join(golden_dir, "extract_full_%s.npz" % shape))
    w, h = (int(v) for v in g["shape"])
    img = synth.synth_image(w, h, seed=int(g["seed"]))
    p = g["params"]
    ex = fe.ORBextractor(int(p[0]), float(p[1]), int(p[2]), int(p[3]), int(p[4]))
    kp, desc = ex(img)
    gk = g["keypoints"]
    assert len(kp) == len(gk)
    for i, f in enumerate(["x", "y", "size", "angle", "response"]):
        assert np.array_equal(kp[f], gk[:, i]), f
    assert np.array_equal(kp["octave"], gk[:, 5].astype(np.int32))
    assert np.array_equal(desc, g["descriptors"])
    pyr = ex.mvImagePyramid
    for lv in range(int(p[2])):
        d = np.frombuffer(hashlib.sha256(np.ascontiguousarray(pyr[lv]).tobytes()).digest(), np.uint8)
        assert np.array_equal(d, g["level_sha256"][lv]), "pyramid level %d" % lv


@pytest.mark.parametrize("shape,seed", [("tum", 0), ("kitti", 0), ("euroc", 0), ("kitti", 3)])
def test_full_size_vs_oracle_with_stage_dumps(fe, oracle, shape, seed):
    s = synth.SETTINGS[shape]
    w, h = synth.SHAPES[shape]
    img = synth.synth_image(w, h, seed)
    params = (s["nfeatures"], s["scale_factor"], s["nlevels"], s["ini_th"], s["min_th"])
    oex = oracle.Extractor(*params)
    okp, odesc = oex.extract(img)
    ex = fe.ORBextractor(*params)
    kp, desc = ex(img)
    h_ = ex._h
    for lv in range(s["nlevels"]):
        assert np.array_equal(h_.level(lv), oex.level(lv)), "pyramid level %d" % lv
        xs, ys, sc = h_.candidates(lv)
        oxs, oys, osc = oex.candidates(lv)
        assert np.array_equal(xs, oxs) and np.array_equal(ys, oys) and np.array_equal(sc, osc), "FAST level %d" % lv
        kx, ky, ks = h_.level_keypoints(lv)
        ok = oex.level_keypoints(lv)
        assert np.array_equal(kx + 16, ok["x"].astype(np.int32)) and np.array_equal(ky + 16, ok["y"].astype(np.int32)), \
            "quadtree level %d" % lv
    _assert_same(kp, desc, okp, odesc)
    assert list(ex.mnFeaturesPerLevel) == list(oex.info.features_per_level)[:s["nlevels"]]
    assert np.array_equal(ex.GetScaleFactors(), np.array(oex.info.scale_factors[:s["nlevels"]], np.float32))


@pytest.mark.parametrize("kind", ["noise", "flat", "gradient"])
def test_stress_images(fe, oracle, kind):
    w, h = 752, 480
    if kind == "noise":
        img = synth.noise_image(w, h, 5)
    elif kind == "flat":
        img = synth.flat_image(w, h)
    else:  # weak texture: exercises the per-cell minThFAST fallback
        rng = np.random.default_rng(9)
        img = np.clip(100 + np.cumsum(rng.integers(-3, 4, (h, w)), axis=1) % 24, 0, 255).astype(np.uint8)
    params = (1200, 1.2, 8, 20, 7)
    okp, odesc = oracle.Extractor(*params).extract(img)
    kp, desc = fe.ORBextractor(*params)(img)
    _assert_same(kp, desc, okp, odesc)
    if kind == "flat":
        assert len(kp) == 0


def test_other_parameters(fe, oracle):
    # monocular initialisation extractor: 2 x nFeatures (src/Tracking.cc:141-142); fewer levels; other thresholds
    img = synth.synth_image(640, 480, 21)
    # scale 1.5 and 2.0 take the generic resize kernel, 1.2 and 1.1 the word-based one (k_pyramid.cu)
    for params in [(2000, 1.2, 8, 20, 7), (500, 1.5, 4, 30, 10), (1000, 1.2, 8, 12, 12), (800, 2.0, 3, 20, 7),
                   (1000, 1.1, 10, 20, 7)]:
        okp, odesc = oracle.Extractor(*params).extract(img)
        kp, desc = fe.ORBextractor(*params)(img)
        _assert_same(kp, desc, okp, odesc)


def test_batch_equals_single_and_empty_input(fe, oracle):
    w, h = 752, 480
    imgs = np.stack([synth.synth_image(w, h, 30 + i) for i in range(5)])
    params = (1200, 1.2, 8, 20, 7)
    ex = fe.ORBextractor(*params, max_batch=5)
    kps, descs, counts = ex.extract_batch(imgs)
    oex = oracle.Extractor(*params)
    for b in range(5):
        okp, odesc = oex.extract(imgs[b])
        _assert_same(kps[b][:counts[b]], descs[b][:counts[b]], okp, odesc)
    kp, desc = ex(np.zeros((0, 0), np.uint8))  # empty image: silent return (src/ORBextractor.cc:1068)
    assert len(kp) == 0 and desc.shape == (0, 32)


def test_repeatable_and_handles_are_independent(fe):
    # two handles driven alternately (left/right extractors of src/Frame.cc:89-92); same input -> same output
    left, right = synth.synth_pair(752, 480, 2)
    a = fe.ORBextractor(1200, 1.2, 8, 20, 7)
    b = fe.ORBextractor(1200, 1.2, 8, 20, 7)
    ka, da = a(left)
    kb, db = b(right)
    ka2, da2 = a(left)
    assert np.array_equal(ka, ka2) and np.array_equal(da, da2)
    kb2, db2 = b(right)
    assert np.array_equal(kb, kb2) and np.array_equal(db, db2)


def test_angle_tolerance_report(fe, oracle):
    # count descriptor bits that differ because of an angle difference: must be zero when the angles are bit-equal
    img = synth.synth_image(1241, 376, 8)
    params = (2000, 1.2, 8, 20, 7)
    okp, odesc = oracle.Extractor(*params).extract(img)
    kp, desc = fe.ORBextractor(*params)(img)
    flips = int(np.unpackbits(desc ^ odesc).sum())
    n_angle = int((kp["angle"] != okp["angle"]).sum())
    print("angle differences: %d, descriptor bit flips: %d" % (n_angle, flips))
    assert n_angle == 0 and flips == 0


@pytest.mark.parametrize("w,h,params", [
    (256, 128, (300, 1.2, 6, 20, 7)),     # width a multiple of 128: pitch == width, TMA boxes / edge words reach the pitch
    (384, 256, (600, 1.2, 8, 20, 7)),     # ... on every second level too
    (131, 97, (200, 1.2, 4, 20, 7)),      # tiny, odd sizes: one or two FAST strips, levels without cells
    (1000, 333, (1500, 1.2, 8, 25, 5)),   # width % 4 == 0 but not % 16: strips start 8 bytes off a TMA-aligned column
    (1283, 721, (2500, 1.2, 8, 20, 7)),   # wider than any config, width % 4 == 3 (two partial right-edge words in the blur)
    (641, 479, (1000, 1.2, 8, 40, 40)),   # iniTh == minTh: no second FAST launch
])
def test_odd_shapes(fe, oracle, w, h, params):
    """Shapes chosen against the staging paths of the kernels (TMA start alignment and out-of-range fill, the blur's
    right-edge CTAs, strips at level borders), on a textured image and on noise."""
    for img in (synth.synth_image(w, h, 77), synth.noise_image(w, h, 78)):
        okp, odesc = oracle.Extractor(*params).extract(img)
        kp, desc = fe.ORBextractor(*params)(img)
        assert len(okp) > 0
        _assert_same(kp, desc, okp, odesc)


@pytest.mark.parametrize("offset,row_pad,img_pad", [(0, 0, 0), (1, 0, 0), (7, 3, 0), (13, 39, 5), (16, 128, 64)])
def test_upload_from_device_memory_at_any_alignment(oracle, offset, row_pad, img_pad):
    """orbfe_extract_upload(images_on_device=1): rows at any byte alignment, padded rows, padded images (the library's
    re-pitch kernel, k_input.cu) — level 0 must equal the images byte for byte, and the features the oracle's."""
    import torch
    from my_orb_slam2_b200 import orbfe
    w, h, B = 333, 77, 3
    params = (150, 1.2, 3, 20, 7)
    imgs = np.stack([synth.synth_image(w, h, 300 + i) for i in range(B)])
    row_stride = w + row_pad
    image_stride = row_stride * h + img_pad
    host = np.full(offset + image_stride * B + 64, 0xEE, np.uint8)  # 0xEE: padding that must never reach level 0
    for b in range(B):
        for y in range(h):
            o = offset + b * image_stride + y * row_stride
            host[o:o + w] = imgs[b, y]
    dev = torch.from_numpy(host).cuda()
    ex = orbfe.Extractor(*params, w, h, max_batch=B)
    ex.upload(dev.data_ptr() + offset, B, image_stride=image_stride, row_stride=row_stride, on_device=True)
    ex.run(B)
    ex.sync()
    for b in range(B):
        assert np.array_equal(ex.level(0, b), imgs[b]), "level 0 of image %d" % b
    cap = ex.capacity
    kps = np.zeros((B, cap), orbfe.KEYPOINT_DTYPE)
    desc = np.zeros((B, cap, 32), np.uint8)
    n = np.zeros(B, np.int32)
    ex.download(B, kps, desc, n)
    for b in range(B):
        okp, odesc = oracle.Extractor(*params).extract(imgs[b])
        assert n[b] == len(okp) and np.array_equal(kps[b][:n[b]], okp) and np.array_equal(desc[b][:n[b]], odesc)
