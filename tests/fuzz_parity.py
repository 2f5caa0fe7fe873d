"""Randomised parity cases for the extractor + stereo matcher: CUDA path (through the C ABI) against the CPU oracle.

Not collected by pytest itself (no test_ prefix): tests/test_gpu_fuzz.py runs a bounded number of seeds, and
`python tests/fuzz_parity.py --first 0 --count 500` runs a long sweep on a GPU box.  Every case draws an image size,
extractor parameters (src/ORBextractor.cc:410-470), an image kind chosen against the kernels' special cases (exact
ties for the 3x3 suppression and the quadtree, saturated pixels, empty cells, corners in the border band) and a
stereo baseline, then compares keypoints, descriptors, mvuRight and mvDepth bit for bit.
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))

KINDS = ("synth", "noise", "checker", "blobs", "saturated", "sparse", "gradient_steps", "smooth_noise")


def _box_blur(a, k):
    c = np.cumsum(np.pad(a.astype(np.float64), ((0, 0), (k, k)), mode="edge"), axis=1)
    a = (c[:, 2 * k:] - c[:, :-2 * k]) / (2 * k)
    c = np.cumsum(np.pad(a, ((k, k), (0, 0)), mode="edge"), axis=0)
    return (c[2 * k:] - c[:-2 * k]) / (2 * k)


def make_image(kind, w, h, rng):
    from my_orb_slam2_b200 import synth
    seed = int(rng.integers(0, 1 << 30))
    if kind == "synth":
        return synth.synth_image(w, h, seed, n_rect=int(rng.integers(5, 600)))
    if kind == "noise":
        return synth.noise_image(w, h, seed)
    if kind == "checker":  # exact score ties everywhere: suppression keeps none of a tied pair, quadtree ties by order
        p = int(rng.integers(3, 24))
        yy, xx = np.mgrid[0:h, 0:w]
        lo, hi = int(rng.integers(0, 100)), int(rng.integers(130, 256))
        return np.where(((yy // p) + (xx // p)) % 2 == 0, lo, hi).astype(np.uint8)
    if kind == "blobs":  # bright / dark discs on a flat ground: corners only at minTh in most cells
        img = np.full((h, w), int(rng.integers(60, 200)), np.float64)
        yy, xx = np.mgrid[0:h, 0:w]
        for _ in range(int(rng.integers(3, 80))):
            cx, cy, r = rng.integers(0, w), rng.integers(0, h), rng.integers(1, 9)
            img[(yy - cy) ** 2 + (xx - cx) ** 2 <= r * r] += rng.uniform(-60, 60)
        return np.clip(np.rint(img), 0, 255).astype(np.uint8)
    if kind == "saturated":  # 0 / 255 only, random rectangles: |diff| = 255 (the SWAR compares' extreme)
        img = np.zeros((h, w), np.uint8)
        for _ in range(int(rng.integers(10, 300))):
            x, y = int(rng.integers(-10, w)), int(rng.integers(-10, h))
            rw, rh = int(rng.integers(2, 50)), int(rng.integers(2, 50))
            img[max(y, 0):max(y + rh, 0), max(x, 0):max(x + rw, 0)] ^= 255
        return img
    if kind == "sparse":  # a flat image with a few isolated pixels / tiny crosses, also inside the 19-px border band
        img = np.full((h, w), 100, np.uint8)
        for _ in range(int(rng.integers(1, 40))):
            x, y = int(rng.integers(0, w)), int(rng.integers(0, h))
            img[y, x] = int(rng.integers(0, 256))
            if rng.random() < 0.5:
                img[max(y - 1, 0):y + 2, x] = img[y, x]
        return img
    if kind == "gradient_steps":  # ramps with steps exactly at / one off the FAST thresholds
        yy, xx = np.mgrid[0:h, 0:w]
        q = int(rng.integers(4, 40))
        img = (xx // q) * int(rng.integers(5, 25)) + (yy // q) * int(rng.integers(5, 25))
        return (img % 256).astype(np.uint8)
    # smooth_noise: low-pass noise, corner strengths near the thresholds
    n = rng.normal(0.0, 1.0, (h, w))
    n = _box_blur(n, int(rng.integers(1, 4)))
    n = n / (n.std() + 1e-9) * rng.uniform(8, 40) + 128
    return np.clip(np.rint(n), 0, 255).astype(np.uint8)


def make_right(left, rng):
    """A right view: the left image shifted by a per-band disparity plus a little noise (true matches at several
    disparities; the band edges give occlusion-like mismatches)."""
    h, w = left.shape
    right = np.empty_like(left)
    nb = int(rng.integers(1, 5))
    edges = np.linspace(0, h, nb + 1).astype(int)
    for i in range(nb):
        d = int(rng.integers(0, max(2, min(w // 3, 90))))
        band = left[edges[i]:edges[i + 1]]
        right[edges[i]:edges[i + 1]] = np.roll(band, -d, axis=1)
    if rng.random() < 0.7:
        noise = rng.integers(-2, 3, left.shape)
        right = np.clip(right.astype(np.int64) + noise, 0, 255).astype(np.uint8)
    return right


def draw_case(seed):
    rng = np.random.default_rng([seed, 777])
    kind = KINDS[int(rng.integers(0, len(KINDS)))]
    nlevels = int(rng.integers(1, 9))
    sf = float(rng.choice([1.1, 1.15, 1.2, 1.2, 1.2, 1.25, 1.3, 1.5, 2.0]))
    # the smallest level must still be wider than tall-ish and hold the 2 x 19 px border (the reference divides by
    # zero / reads outside otherwise; the product refuses such sizes with ORBFE_ERR_ARG)
    min_dim = int(np.ceil(int(rng.choice([36, 48, 70, 70, 100])) * sf ** (nlevels - 1)))
    w = int(rng.integers(min_dim, max(min_dim + 1, 1400)))
    h = int(rng.integers(min_dim, max(min_dim + 1, min(w + 1, 900))))
    if rng.random() < 0.25:
        w = (w + 127) // 128 * 128  # pitch == width
    ini = int(rng.integers(5, 60))
    mn = int(rng.integers(1, ini + 1))
    nfeat = int(rng.choice([30, 100, 300, 500, 1000, 1200, 2000, 3000, 5000]))
    fx = float(rng.uniform(200, 900))
    bf = float(fx * rng.uniform(0.05, 0.6))
    return dict(seed=seed, kind=kind, w=w, h=h, params=(nfeat, sf, nlevels, ini, mn), fx=fx, bf=bf, rng=rng)


def run_case(seed, oracle, frontend, orbfe, extractors=None):
    """Returns (description, status) with status in {"ok", "refused", "DIFF: ..."}."""
    c = draw_case(seed)
    rng = c["rng"]
    desc = "seed %d %s %dx%d params %s" % (seed, c["kind"], c["w"], c["h"], c["params"])
    left = make_image(c["kind"], c["w"], c["h"], rng)
    right = make_right(left, rng)
    try:
        eL, eR = frontend.ORBextractor(*c["params"]), frontend.ORBextractor(*c["params"])
        f = frontend.Frame(left, right, eL, eR, c["bf"], c["fx"])
    except orbfe.OrbfeError as e:
        if e.code == orbfe.ERR_ARG:
            return desc, "refused: " + str(e)
        raise
    oL, oR = oracle.Extractor(*c["params"]), oracle.Extractor(*c["params"])
    kl, dl = oL.extract(left)
    kr, dr = oR.extract(right)
    desc += " n=%d/%d" % (len(kl), len(kr))
    for name, a, b in (("left keypoints", f.mvKeys, kl), ("left descriptors", f.mDescriptors, dl),
                       ("right keypoints", f.mvKeysRight, kr), ("right descriptors", f.mDescriptorsRight, dr)):
        if len(a) != len(b) or not np.array_equal(a, b):
            return desc, "DIFF: %s (%d vs %d)" % (name, len(a), len(b))
    if len(kl):
        mb = np.float32(c["bf"]) / np.float32(c["fx"])
        our, odp, on = oracle.stereo_matches(oL, oR, kl, dl, kr, dr, np.float32(c["bf"]), mb)
        if not (np.array_equal(f.mvuRight, our) and np.array_equal(f.mvDepth, odp)):
            bad = int((f.mvuRight != our).sum() + (f.mvDepth != odp).sum())
            return desc, "DIFF: stereo (%d entries)" % bad
        desc += " matched=%d" % on
    return desc, "ok"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--first", type=int, default=0)
    ap.add_argument("--count", type=int, default=100)
    args = ap.parse_args()
    import pyoracle as oracle
    from my_orb_slam2_b200 import frontend, orbfe
    n_ok = n_ref = 0
    bad = []
    for seed in range(args.first, args.first + args.count):
        d, st = run_case(seed, oracle, frontend, orbfe)
        if st == "ok":
            n_ok += 1
        elif st.startswith("refused"):
            n_ref += 1
            print(d, st, flush=True)
        else:
            bad.append((d, st))
            print("FAIL", d, st, flush=True)
    print("fuzz: %d ok, %d refused (ORBFE_ERR_ARG), %d different" % (n_ok, n_ref, len(bad)))
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
