"""GPU parity of the DBoW2 vocabulary transform against the CPU oracle: word / node of every
descriptor, the BowVector (ids exact, values BIT-equal doubles: same additions in the same order) and the
FeatureVector, for every scoring / weighting combination the reference's enums allow; the text loader; the
device-resident variant chained behind the extractor; and SearchByBoW fed with the GPU's feature vectors."""
import numpy as np
import pytest

from cases import extract_with_oracle, views
from my_orb_slam2_b200 import synth

pytestmark = pytest.mark.gpu


def _both(oracle, voc, scoring, weighting):
    from my_orb_slam2_b200 import orbfe
    args = (voc["k"], voc["L"], scoring, weighting, voc["parent"], voc["is_leaf"], voc["desc"], voc["weight"])
    return oracle.Vocabulary.from_arrays(*args), orbfe.Vocabulary.from_arrays(*args)


def _same(got, want):
    names = ["bow_ids", "bow_vals", "fv_nodes", "fv_offsets", "fv_features", "word_of", "node_of", "weight_of"]
    for name, g, w in zip(names, got, want):
        if g is None:
            continue
        assert g.dtype == w.dtype and np.array_equal(g, w), name


@pytest.mark.parametrize("k,L,scoring,weighting,levelsup,n", [
    (10, 4, 0, 0, 2, 2000), (10, 4, 0, 0, 4, 2000), (10, 4, 1, 0, 2, 1500), (10, 4, 5, 0, 2, 1500),
    (10, 4, 5, 1, 1, 700), (10, 4, 2, 2, 2, 1500), (10, 4, 0, 3, 3, 1500), (3, 6, 0, 0, 4, 1000),
    (20, 2, 0, 0, 1, 3000), (10, 4, 0, 0, 2, 1), (10, 4, 0, 0, 2, 33), (10, 5, 0, 0, 3, 9000)])
def test_transform_vs_oracle(oracle, k, L, scoring, weighting, levelsup, n):
    voc = synth.vocabulary_case(k, L, seed=k + L)
    ov, gv = _both(oracle, voc, scoring, weighting)
    assert (gv.k, gv.L, gv.scoring, gv.weighting, gv.n_nodes, gv.n_words) == \
           (ov.k, ov.L, ov.scoring, ov.weighting, ov.n_nodes, ov.n_words)
    desc = synth.descriptors_near_words(voc, n, seed=n)
    want = ov.transform(desc, levelsup)
    got = gv.transform(desc, levelsup)
    _same(got, want)
    if n >= 700:
        assert len(want[0]) > 100 and (want[7] <= 0).sum() > 0


def test_transform_edge_cases(oracle):
    voc = synth.vocabulary_case(10, 3, seed=5, stop_frac=1.0)  # every word stopped: empty containers
    ov, gv = _both(oracle, voc, 0, 0)
    desc = synth.descriptors_near_words(voc, 100, seed=1)
    want, got = ov.transform(desc, 1), gv.transform(desc, 1)
    assert len(want[0]) == 0 and len(want[2]) == 0
    _same(got, want)
    got = gv.transform(desc[:0], 1)
    assert len(got[0]) == 0 and len(got[2]) == 0
    with pytest.raises(Exception):
        gv.transform(np.zeros((20000, 32), np.uint8), 1)  # above the documented capacity: loud, not wrong


def test_text_loader(oracle, tmp_path):
    from my_orb_slam2_b200 import orbfe
    voc = synth.vocabulary_case(10, 3, seed=9)
    path = tmp_path / "voc.txt"
    path.write_text(synth.vocabulary_text(voc, 0, 0))
    gv = orbfe.Vocabulary.load_text(str(path))
    ov = oracle.Vocabulary.load_text(str(path))
    assert (gv.k, gv.L, gv.n_nodes, gv.n_words) == (ov.k, ov.L, ov.n_nodes, ov.n_words) == (10, 3, len(voc["parent"]),
                                                                                           int(voc["is_leaf"].sum()))
    desc = synth.descriptors_near_words(voc, 1200, seed=3)
    _same(gv.transform(desc, 2), ov.transform(desc, 2))
    (tmp_path / "bad.txt").write_text("99 3 0 0\n")
    with pytest.raises(Exception):
        orbfe.Vocabulary.load_text(str(tmp_path / "bad.txt"))


def test_compute_bow_chain(oracle):
    """Frame::ComputeBoW on the extractor's device-resident descriptors, then SearchByBoW with those feature vectors."""
    import ctypes as C
    from my_orb_slam2_b200 import frontend as fe, orbfe
    params, img, oex, (okp, odesc) = extract_with_oracle(oracle, "kitti", 0)
    ex = fe.ORBextractor(*params)
    kp, desc = ex(img)
    assert np.array_equal(desc, odesc)
    voc = synth.vocabulary_case(10, 4, seed=1)
    ov, gv = _both(oracle, voc, 0, 0)
    h = ex.handle(img.shape[1], img.shape[0])
    dk, dd, dc, cap = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_int()
    orbfe.check(orbfe.lib().orbfe_extractor_device_results(h._h, C.byref(dk), C.byref(dd), C.byref(dc), C.byref(cap)))
    got = gv.transform(None, 4, device_ptr=dd.value, n=len(kp))
    want = ov.transform(odesc, 4)
    _same(got, want)
    # the feature vectors feed SearchByBoW (KeyFrame vs Frame): second view = same features, flipped bits
    rng = np.random.default_rng(3)
    d2 = synth.flip_bits(odesc, 0.03, rng)
    fv2 = gv.transform(d2, 4)
    sf = synth.scale_factors(8)

    def kfv(k, d, fv, has):
        return dict(keys_un=k, u_right=np.full(len(k), -1.0, np.float32), desc=d, has_mp=has, node_ids=fv[2],
                    node_offsets=fv[3], node_features=fv[4], scale_factors=sf, level_sigma2=sf * sf)

    has = (rng.random(len(kp)) < 0.8).astype(np.uint8)
    k1, k2 = kfv(okp, odesc, got, has), kfv(okp, d2, fv2, has)
    wm, wn = oracle.search_by_bow(views(oracle, kf=k1), views(oracle, kf=k2), False, 0.7, True)
    gm, gn = fe.ORBmatcher(0.7, True).SearchByBoW(views(fe, kf=k1), views(fe, kf=k2), False)
    assert gn == wn and wn > 500 and np.array_equal(gm, wm)
