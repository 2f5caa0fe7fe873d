"""The data-parallel quadtree algorithm of k_quadtree.cu (numpy model, tests/quadtree_model.py) must give the
oracle's DistributeOctTree result (src/ORBextractor.cc:539-765): same kept keypoints, same list order."""
import numpy as np
import pytest

from quadtree_model import distribute


def _cands_from_extractor(oracle, img, params):
    ex = oracle.Extractor(*params)
    ex.extract(img)
    return ex


@pytest.mark.parametrize("seed,w,h,nf", [(1, 320, 240, 300), (2, 400, 200, 500), (3, 640, 200, 1000), (4, 200, 160, 60)])
def test_model_matches_oracle_on_images(oracle, seed, w, h, nf):
    from my_orb_slam2_b200 import synth
    img = synth.synth_image(w, h, seed=seed, n_rect=int(w * h / 600)) if seed != 4 else synth.noise_image(w, h, seed)
    ex = _cands_from_extractor(oracle, img, (nf, 1.2, 6, 20, 7))
    for lv in range(6):
        xs, ys, sc = ex.candidates(lv)
        lvl = ex.level(lv)
        lh, lw = lvl.shape
        width, height = lw - 32, lh - 32
        if width < 30 or height < 30:
            continue
        n_cols, n_rows = int(np.float32(width) / np.float32(30)), int(np.float32(height) / np.float32(30))
        w_cell, h_cell = int(np.ceil(np.float32(width) / n_cols)), int(np.ceil(np.float32(height) / n_rows))
        quota = ex.info.features_per_level[lv]
        want = oracle.distribute_octtree(xs, ys, sc, 16, lw - 16, 16, lh - 16, quota)
        got = distribute(xs, ys, sc, width, height, quota, w_cell, h_cell)
        assert list(want) == got, "level %d" % lv


def test_model_random_ties(oracle):
    rng = np.random.default_rng(0)
    for trial in range(30):
        width, height = int(rng.integers(60, 700)), int(rng.integers(40, 200))
        if width < height:
            width, height = height * 2, width
        n = int(rng.integers(1, 1500))
        pts = rng.permutation(width * height)[:n]
        xs = (pts % width).astype(np.int32)
        ys = (pts // width).astype(np.int32)
        # the FAST candidates come in cell order: emulate with a fixed cell size and sort
        w_cell, h_cell = 31, 32
        xs += 3
        ys += 3
        order = np.lexsort((xs, ys, (xs - 3) // w_cell, (ys - 3) // h_cell))
        xs, ys = xs[order], ys[order]
        sc = rng.integers(7, 40, n).astype(np.int32)  # few distinct scores -> many ties
        quota = int(rng.integers(1, 400))
        want = oracle.distribute_octtree(xs, ys, sc, 16, 16 + width + 6, 16, 16 + height + 6, quota)
        got = distribute(xs, ys, sc, width + 6, height + 6, quota, w_cell, h_cell)
        assert list(want) == got, "trial %d" % trial
