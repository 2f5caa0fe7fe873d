"""Parity on BASELINE.json's configs AS CONFIGURED: the KITTI batch of 64 stereo pairs the bench
times (configs[1]), the EuRoC batch of 256 stereo pairs on one GPU (configs[2]) and the 8192 x 1M brute force
(configs[4]) against the CPU oracle, bit-exact.  The oracle legs are sized to finish in seconds: every pair of the KITTI
batch, a spread sample of the EuRoC batch (all 256 are compared with the GPU's own single-pair result), 256 sampled
queries of the brute force against the full 1M map."""
import numpy as np
import pytest

from my_orb_slam2_b200 import synth

pytestmark = pytest.mark.gpu


def _params(s):
    return (s["nfeatures"], s["scale_factor"], s["nlevels"], s["ini_th"], s["min_th"])


def _oracle_pair(oracle, s, left, right):
    p = _params(s)
    oL, oR = oracle.Extractor(*p), oracle.Extractor(*p)
    kl, dl = oL.extract(left)
    kr, dr = oR.extract(right)
    mbf = np.float32(s["bf"])
    ur, dp, n = oracle.stereo_matches(oL, oR, kl, dl, kr, dr, mbf, mbf / np.float32(s["fx"]))
    return kl, dl, kr, dr, ur, dp, n


def _run_batch(shape, n_pairs, n_unique):
    from my_orb_slam2_b200 import orbfe
    s = synth.SETTINGS[shape]
    w, h = synth.SHAPES[shape]
    uniq = [synth.synth_pair(w, h, seed=i) for i in range(n_unique)]
    lefts = np.stack([uniq[i % n_unique][0] for i in range(n_pairs)])
    rights = np.stack([uniq[i % n_unique][1] for i in range(n_pairs)])
    L = orbfe.Extractor(*_params(s), w, h, max_batch=n_pairs)
    R = orbfe.Extractor(*_params(s), w, h, max_batch=n_pairs)
    mbf = np.float32(s["bf"])
    out = orbfe.stereo_frame(L, R, lefts, rights, mbf, mbf / np.float32(s["fx"]))
    return s, lefts, rights, out


def _check_pair(oracle, s, lefts, rights, out, b):
    kl, dl, nl, kr, dr, nr, ur, dp, nm = out
    okl, odl, okr, odr, our, odp, on = _oracle_pair(oracle, s, lefts[b], rights[b])
    assert nl[b] == len(okl) and nr[b] == len(okr), "pair %d" % b
    assert np.array_equal(kl[b][:nl[b]], okl) and np.array_equal(dl[b][:nl[b]], odl), "left of pair %d" % b
    assert np.array_equal(kr[b][:nr[b]], okr) and np.array_equal(dr[b][:nr[b]], odr), "right of pair %d" % b
    assert np.array_equal(ur[b][:nl[b]], our) and np.array_equal(dp[b][:nl[b]], odp), "stereo of pair %d" % b
    assert nm[b] == on


def test_kitti_batch_64_as_benched(oracle):
    """configs[1] as bench.py runs it: 64 distinct KITTI-shaped pairs in one batched call; every pair against the oracle."""
    s, lefts, rights, out = _run_batch("kitti", 64, 64)
    for b in range(64):
        _check_pair(oracle, s, lefts, rights, out, b)


def test_euroc_batch_256_one_gpu(oracle):
    """configs[2] on one GPU: one batched call over 256 EuRoC-shaped pairs (32 distinct pairs, cycled, so that image
    slot b and slot b+32 must agree too); 16 pairs spread over the batch against the oracle."""
    s, lefts, rights, out = _run_batch("euroc", 256, 32)
    kl, dl, nl, kr, dr, nr, ur, dp, nm = out
    for b in range(32, 256):
        a = b % 32
        assert nl[b] == nl[a] and nr[b] == nr[a] and nm[b] == nm[a]
        assert np.array_equal(kl[b][:nl[b]], kl[a][:nl[a]]) and np.array_equal(dl[b][:nl[b]], dl[a][:nl[a]])
        assert np.array_equal(kr[b][:nr[b]], kr[a][:nr[a]]) and np.array_equal(dr[b][:nr[b]], dr[a][:nr[a]])
        assert np.array_equal(ur[b][:nl[b]], ur[a][:nl[a]]) and np.array_equal(dp[b][:nl[b]], dp[a][:nl[a]])
    for b in list(range(0, 256, 17)) + [255]:
        _check_pair(oracle, s, lefts, rights, out, b)


def test_bruteforce_1m_sampled_queries_vs_oracle(oracle):
    """configs[4] on one GPU: 8192 queries x 1 048 576 map descriptors; 256 queries spread over the set are checked
    against the oracle's double loop over the FULL map (best and second keys, ties to the lowest index)."""
    from my_orb_slam2_b200 import orbfe
    q, m = synth.bruteforce_case(8192, 1 << 20, 4096, seed=1)
    m[777777] = m[5]          # exact duplicates far apart: the lower index must win, the other must be second
    q[0] = m[5]
    best, second = orbfe.Matcher().hamming_bruteforce(q, m, 0)
    sample = np.unique(np.concatenate([[0, 1, 8191], np.arange(0, 8192, 32)]))[:256]
    ob, os_ = oracle.hamming_bruteforce(np.ascontiguousarray(q[sample]), m, 0)
    assert np.array_equal(best[sample], ob) and np.array_equal(second[sample], os_)
    assert int(best[0]) == 5 and int(second[0]) == 777777
