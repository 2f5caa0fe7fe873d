"""Seeded synthetic inputs of TUM / KITTI / EuRoC shape (no datasets offline).

SURVEY.md 8(d): mid-grey + low-frequency background + filled rectangles with
random intensity steps + pixel noise, which gives a realistic FAST corner
density (~10 k candidates over 8 levels at KITTI shape).  Stereo pairs render
the same rectangle list twice, each rectangle shifted left in the right image
by its own (quarter-pixel) disparity, so Frame::ComputeStereoMatches has true
matches with sub-pixel offsets.  numpy only, deterministic for a given seed.
"""
import numpy as np

SHAPES = {
    "tum": (640, 480),     # Examples/Monocular/TUM1.yaml
    "kitti": (1241, 376),  # Examples/Stereo/KITTI00-02.yaml
    "euroc": (752, 480),   # Examples/Stereo/EuRoC.yaml
}

# (nFeatures, scaleFactor, nLevels, iniThFAST, minThFAST, fx, bf) from the reference's settings files
SETTINGS = {
    "tum": dict(nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, fx=517.306408, bf=0.0),
    "kitti": dict(nfeatures=2000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, fx=718.856, bf=386.1448),
    "euroc": dict(nfeatures=1200, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, fx=435.2046959714599,
                  bf=47.90639384423901),
}

_SS = 4  # horizontal supersampling (quarter-pixel disparities)


def _upsample_bilinear(small, h, w):
    sh, sw = small.shape
    ys = np.linspace(0, sh - 1, h)
    xs = np.linspace(0, sw - 1, w)
    y0 = np.floor(ys).astype(np.int64)
    x0 = np.floor(xs).astype(np.int64)
    y1 = np.minimum(y0 + 1, sh - 1)
    x1 = np.minimum(x0 + 1, sw - 1)
    fy = (ys - y0)[:, None]
    fx = (xs - x0)[None, :]
    top = small[y0][:, x0] * (1 - fx) + small[y0][:, x1] * fx
    bot = small[y1][:, x0] * (1 - fx) + small[y1][:, x1] * fx
    return top * (1 - fy) + bot * fy


def _scene(w, h, seed, n_rect=None):
    rng = np.random.default_rng(seed)
    if n_rect is None:
        n_rect = int(round(400 * (w * h) / (1241.0 * 376.0)))
    pad = 128  # background is rendered wider so that the right view can be a shifted crop
    bg_small = rng.normal(0.0, 1.0, (h // 6 + 2, (w + pad) // 6 + 2))
    bg = 110.0 + 12.0 * _upsample_bilinear(bg_small, h, w + pad)
    rects = []
    for _ in range(n_rect):
        rw = int(rng.integers(6, 61))
        rh = int(rng.integers(6, 61))
        x = int(rng.integers(-20, w + 20))
        y = int(rng.integers(-20, h + 20))
        step = float(rng.uniform(-90, 90))
        disp4 = int(rng.integers(2 * _SS, 96 * _SS + 1))  # disparity in quarter pixels
        rects.append((disp4, x, y, rw, rh, step))
    rects.sort(key=lambda r: r[0])  # far first, near rectangles occlude
    return bg, rects, pad


def _render(w, h, bg, rects, pad, right, noise_rng):
    bgd = 2 if right else 0
    base = bg[:, bgd:bgd + w]
    canvas = np.repeat(base, _SS, axis=1)  # h x (w*SS)
    for disp4, x, y, rw, rh, step in rects:
        x0 = x * _SS - (disp4 if right else 0)
        x1 = x0 + rw * _SS
        y0, y1 = max(y, 0), min(y + rh, h)
        x0c, x1c = max(x0, 0), min(x1, w * _SS)
        if y1 <= y0 or x1c <= x0c:
            continue
        canvas[y0:y1, x0c:x1c] = 110.0 + step
    img = canvas.reshape(h, w, _SS).mean(axis=2)
    img = img + noise_rng.normal(0.0, 2.0, (h, w))
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def synth_image(w, h, seed=0, n_rect=None):
    """One grayscale frame (H x W uint8)."""
    bg, rects, pad = _scene(w, h, seed, n_rect)
    return _render(w, h, bg, rects, pad, False, np.random.default_rng([seed, 1]))


def synth_pair(w, h, seed=0, n_rect=None):
    """Rectified stereo pair (left, right), each H x W uint8."""
    bg, rects, pad = _scene(w, h, seed, n_rect)
    left = _render(w, h, bg, rects, pad, False, np.random.default_rng([seed, 1]))
    right = _render(w, h, bg, rects, pad, True, np.random.default_rng([seed, 2]))
    return left, right


def noise_image(w, h, seed=0):
    """Stress input: uniform noise (every pixel neighbourhood is a candidate region)."""
    return np.random.default_rng([seed, 3]).integers(0, 256, (h, w), dtype=np.uint8)


def flat_image(w, h, value=128):
    """Stress input: no keypoints at all (thr-7 fallback in every cell, nkeypoints == 0 branch)."""
    return np.full((h, w), value, np.uint8)


def random_descriptors(n, seed):
    return np.random.default_rng([seed, 4]).integers(0, 256, (n, 32), dtype=np.uint8)


def flip_bits(desc, p, rng):
    """Flip each of the 256 bits of every descriptor with probability p."""
    flips = rng.random((desc.shape[0], 256)) < p
    return desc ^ np.packbits(flips, axis=1, bitorder="little")
