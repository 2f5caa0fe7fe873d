"""CPU checks of the oracle's DBoW2 transform against an independent pure-Python reading of
TemplatedVocabulary.h:1127-1262 (dict instead of std::map, sorted at the end) on small trees, and of the text loader."""
import math

import numpy as np
import pytest

from my_orb_slam2_b200 import synth


def py_transform(voc, desc, levelsup, scoring=0, weighting=0):
    parent, weight, ndesc = voc["parent"], voc["weight"], voc["desc"]
    n = len(parent)
    children = [[] for _ in range(n)]
    word_id, nw = np.zeros(n, np.int64), 0
    for i in range(1, n):
        children[parent[i]].append(i)
        if voc["is_leaf"][i]:
            word_id[i] = nw
            nw += 1
    bow, fv = {}, {}
    nid_level = voc["L"] - levelsup
    for i, d in enumerate(desc):
        final, nid, level = 0, 0, 0
        while True:
            level += 1
            ch = children[final]
            dist = [int(np.unpackbits(d ^ ndesc[c]).sum()) for c in ch]
            final = ch[int(np.argmin(dist))]  # argmin returns the first minimum
            if level == nid_level:
                nid = final
            if not children[final]:
                break
        w = float(weight[final])
        if w > 0:
            wid = int(word_id[final])
            if weighting in (0, 1):
                bow[wid] = bow[wid] + w if wid in bow else w
            elif wid not in bow:
                bow[wid] = w
            fv.setdefault(nid, []).append(i)
    ids = sorted(bow)
    vals = [bow[k] for k in ids]
    must = scoring != 5
    if weighting in (0, 1) and ids and not must:
        vals = [v / float(len(ids)) for v in vals]
    if must:
        norm = 0.0
        for v in vals:
            norm += abs(v) if scoring != 1 else v * v
        if scoring == 1:
            norm = math.sqrt(norm)
        if norm > 0:
            vals = [v / norm for v in vals]
    return ids, vals, fv


@pytest.mark.parametrize("k,L,scoring,weighting,levelsup", [(4, 3, 0, 0, 1), (10, 3, 0, 0, 2), (3, 4, 1, 1, 2),
                                                            (5, 3, 5, 0, 1), (5, 3, 0, 2, 4), (6, 2, 2, 3, 1)])
def test_transform_vs_python(oracle, k, L, scoring, weighting, levelsup):
    voc = synth.vocabulary_case(k, L, seed=k * 10 + L, prune=0.15, stop_frac=0.15)
    desc = synth.descriptors_near_words(voc, 300, seed=1)
    ov = oracle.Vocabulary.from_arrays(k, L, scoring, weighting, voc["parent"], voc["is_leaf"], voc["desc"], voc["weight"])
    ids, vals, fvn, fvo, fvf, word, node, wt = ov.transform(desc, levelsup)
    pids, pvals, pfv = py_transform(voc, desc, levelsup, scoring, weighting)
    assert list(ids) == pids and len(pids) > 10
    assert np.array_equal(vals, np.array(pvals, np.float64))  # same additions in the same order: bit-equal doubles
    assert list(fvn) == sorted(pfv)
    for j, nd in enumerate(fvn):
        assert list(fvf[fvo[j]:fvo[j + 1]]) == pfv[int(nd)]
    if scoring in (0, 2, 3, 4):
        assert abs(vals.sum() - 1.0) < 1e-12
    assert (wt <= 0).sum() > 0  # some features hit stopped words


def test_text_loader_round_trip(oracle, tmp_path):
    voc = synth.vocabulary_case(5, 3, seed=7)
    path = tmp_path / "voc.txt"
    path.write_text(synth.vocabulary_text(voc, 0, 0))
    a = oracle.Vocabulary.load_text(str(path))
    b = oracle.Vocabulary.from_arrays(5, 3, 0, 0, voc["parent"], voc["is_leaf"], voc["desc"], voc["weight"])
    assert (a.k, a.L, a.scoring, a.weighting, a.n_nodes, a.n_words) == (b.k, b.L, b.scoring, b.weighting, b.n_nodes, b.n_words)
    desc = synth.descriptors_near_words(voc, 200, seed=2)
    for x, y in zip(a.transform(desc, 2), b.transform(desc, 2)):
        assert np.array_equal(x, y)
