"""CPU oracle vs the committed cv2-4.13 golden fixtures (tests/golden/gen_golden.py) and known-answer vectors.

These pin the oracle: every pixel primitive in the fixtures came from the real OpenCV call the reference makes
(src/ORBextractor.cc:103,830,835,1108,1142)."""
import os

import numpy as np
import pytest

CASES = ["small_a", "small_b", "small_noise"]


def _extractor(oracle, params, flags=0):
    p = params
    return oracle.Extractor(int(p[0]), float(p[1]), int(p[2]), int(p[3]), int(p[4]), flags)


def test_primitives_match_cv2_fixture(oracle, golden_dir):
    g = np.load(os.path.join(golden_dir, "primitives.npz"))
    assert np.array_equal(oracle.resize_linear(g["src"], 109, 81), g["resized"])
    assert np.array_equal(oracle.gaussian_blur7(g["src"]), g["blurred"])
    at = np.array([oracle.fast_atan2(y, x) for y, x in zip(g["atan_y"], g["atan_x"])], np.float32)
    assert np.array_equal(at, g["atan"])
    xs, ys, sc = oracle.fast9_16(g["fast_roi"], 20, True)
    assert np.array_equal(np.stack([xs, ys, sc], 1).reshape(-1, 3), g["fast_kps"].reshape(-1, 3))


@pytest.mark.parametrize("name", CASES)
def test_extract_matches_cv2_fixture(oracle, golden_dir, name):
    g = np.load(os.path.join(golden_dir, "extract_%s.npz" % name))
    ex = _extractor(oracle, g["params"])
    kp, desc = ex.extract(g["image"])
    gk = g["keypoints"]
    assert len(kp) == len(gk)
    assert np.array_equal(kp["x"], gk[:, 0]) and np.array_equal(kp["y"], gk[:, 1])
    assert np.array_equal(kp["size"], gk[:, 2])
    assert np.array_equal(kp["angle"], gk[:, 3])  # bit-equal, tolerance of north_star is 1e-3 deg
    assert np.array_equal(kp["response"], gk[:, 4])
    assert np.array_equal(kp["octave"], gk[:, 5].astype(np.int32))
    assert np.array_equal(desc, g["descriptors"])
    nl = int(g["params"][2])
    assert np.array_equal(ex.level(nl - 1), g["level_last"])
    for lv in range(nl):
        xs, ys, sc = ex.candidates(lv)
        assert np.array_equal(np.stack([xs, ys, sc], 1).reshape(-1, 3), g["cand_%d" % lv].reshape(-1, 3))


def _sha(a):
    import hashlib
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), np.uint8)


@pytest.mark.parametrize("shape", ["tum", "kitti", "euroc"])
def test_full_size_matches_cv2_fixture(oracle, golden_dir, shape):
    """BASELINE.json's three shapes at full size and full feature count against the cv2 4.13 chain: every pyramid level
    (the 7 resize transitions), every level's GaussianBlur, every level's candidate list, keypoints and descriptors."""
    from my_orb_slam2_b200 import synth
    g = np.load(os.path.join(golden_dir, "extract_full_%s.npz" % shape))
    w, h = (int(v) for v in g["shape"])
    img = synth.synth_image(w, h, seed=int(g["seed"]))
    assert np.array_equal(_sha(img), g["image_sha256"]), "the synthetic generator changed: regenerate tests/golden"
    ex = _extractor(oracle, g["params"])
    kp, desc = ex.extract(img)
    gk = g["keypoints"]
    assert len(kp) == len(gk)
    for i, f in enumerate(["x", "y", "size", "angle", "response"]):
        assert np.array_equal(kp[f], gk[:, i]), f
    assert np.array_equal(kp["octave"], gk[:, 5].astype(np.int32))
    assert np.array_equal(desc, g["descriptors"])
    nl = int(g["params"][2])
    assert np.array_equal(ex.level(nl - 1), g["level_last"])
    for lv in range(nl):
        lvl = ex.level(lv)
        assert [lvl.shape[1], lvl.shape[0]] == list(g["level_dims"][lv])
        assert np.array_equal(_sha(lvl), g["level_sha256"][lv]), "pyramid level %d" % lv
        assert np.array_equal(_sha(oracle.gaussian_blur7(lvl)), g["blur_sha256"][lv]), "blur of level %d" % lv
        xs, ys, sc = ex.candidates(lv)
        c = np.stack([xs, ys, sc], 1).reshape(-1, 3).astype(np.int32)
        assert len(c) == int(g["cand_counts"][lv])
        assert np.array_equal(_sha(c), g["cand_sha256"][lv]), "candidates of level %d" % lv


def test_quotas_and_umax(oracle):
    # quotas and umax of the constructor (src/ORBextractor.cc:435-469)
    want = {1000: [217, 181, 151, 126, 105, 87, 73, 60], 2000: [434, 362, 302, 251, 209, 175, 145, 122],
            1200: [261, 217, 181, 151, 126, 105, 87, 72]}
    for nf, q in want.items():
        ex = oracle.Extractor(nf, 1.2, 8, 20, 7)
        assert list(ex.info.features_per_level)[:8] == q
        assert list(ex.info.umax) == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]


def test_descriptor_distance_kats(oracle):
    z = np.zeros(32, np.uint8)
    f = np.full(32, 255, np.uint8)
    assert oracle.descriptor_distance(z, z) == 0
    assert oracle.descriptor_distance(z, f) == 256
    rng = np.random.default_rng(3)
    for _ in range(200):
        a = rng.integers(0, 256, 32, dtype=np.uint8)
        b = rng.integers(0, 256, 32, dtype=np.uint8)
        assert oracle.descriptor_distance(a, b) == int(np.unpackbits(a ^ b).sum())
    one = z.copy()
    one[17] = 0x10
    assert oracle.descriptor_distance(z, one) == 1


def test_sincos_matches_correctly_rounded(oracle):
    import math
    rng = np.random.default_rng(5)
    bad = 0
    for a in rng.uniform(0, 360, 20000).astype(np.float32):
        ang = np.float32(a) * np.float32(np.pi / 180.0)
        c, s = oracle.sincos_deg(a)
        bad += (c != np.float32(math.cos(float(ang)))) + (s != np.float32(math.sin(float(ang))))
    assert bad == 0


def test_flat_and_empty(oracle):
    from my_orb_slam2_b200 import synth
    ex = oracle.Extractor(500, 1.2, 8, 20, 7)
    kp, desc = ex.extract(synth.flat_image(320, 240))
    assert len(kp) == 0 and desc.shape == (0, 32)


def test_hamming_bruteforce_ties_lowest_index(oracle):
    q = np.zeros((2, 32), np.uint8)
    m = np.zeros((5, 32), np.uint8)
    m[0, 0] = 3
    m[1, 0] = 1
    m[2, 0] = 1
    best, second = oracle.hamming_bruteforce(q, m, index_base=100)
    assert int(best[0]) == (0 << 32) | 103 and int(second[0]) == (0 << 32) | 104


def test_input_side_matches_cv2_fixture(oracle, golden_dir):
    """cvtColor *2GRAY and remap INTER_LINEAR against cv2 4.13 outputs (tests/golden/gen_golden_input.py)."""
    import os
    g = np.load(os.path.join(golden_dir, "input_side.npz"))
    for name, rgb in (("rgb", True), ("bgr", False), ("rgba", True), ("bgra", False)):
        assert np.array_equal(oracle.cvt_gray(g["color_" + name], rgb), g["gray_" + name]), name
    for i in range(2):
        got = oracle.remap_linear(g["remap_src_%d" % i], g["remap_mx_%d" % i], g["remap_my_%d" % i])
        assert np.array_equal(got, g["remap_dst_%d" % i]), i
    assert (g["remap_dst_1"][:10] == 0).mean() > 0.1  # the constant border was exercised
