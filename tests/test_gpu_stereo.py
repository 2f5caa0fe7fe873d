"""GPU parity of Frame::ComputeStereoMatches (src/Frame.cc:496-686) against the CPU oracle: mvuRight / mvDepth
bit-exact (they are fp32 results of explicitly rounded operations), match count equal."""
import numpy as np
import pytest

from my_orb_slam2_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape,seed", [("kitti", 0), ("euroc", 0), ("euroc", 1)])
def test_stereo_matches_vs_oracle(oracle, shape, seed):
    from my_orb_slam2_b200 import frontend
    s = synth.SETTINGS[shape]
    w, h = synth.SHAPES[shape]
    left, right = synth.synth_pair(w, h, seed)
    params = (s["nfeatures"], s["scale_factor"], s["nlevels"], s["ini_th"], s["min_th"])
    oL, oR = oracle.Extractor(*params), oracle.Extractor(*params)
    kl, dl = oL.extract(left)
    kr, dr = oR.extract(right)
    mb = np.float32(s["bf"]) / np.float32(s["fx"])
    our, odp, on = oracle.stereo_matches(oL, oR, kl, dl, kr, dr, np.float32(s["bf"]), mb)
    f = frontend.Frame(left, right, frontend.ORBextractor(*params), frontend.ORBextractor(*params), s["bf"], s["fx"])
    assert f.N == len(kl) and np.array_equal(f.mvKeys, kl) and np.array_equal(f.mvKeysRight, kr)
    assert np.array_equal(f.mvuRight, our)
    assert np.array_equal(f.mvDepth, odp)
    assert int((f.mvuRight >= 0).sum()) == on and on > 50


def test_stereo_batch_and_no_match_cases(oracle):
    from my_orb_slam2_b200 import orbfe
    w, h = 752, 480
    s = synth.SETTINGS["euroc"]
    params = (s["nfeatures"], s["scale_factor"], s["nlevels"], s["ini_th"], s["min_th"])
    pairs = [synth.synth_pair(w, h, 40 + i) for i in range(3)]
    # image 3: right view is unrelated noise -> (almost) no matches; image 4: flat -> zero keypoints
    lefts = np.stack([p[0] for p in pairs] + [synth.synth_image(w, h, 50), synth.flat_image(w, h)])
    rights = np.stack([p[1] for p in pairs] + [synth.noise_image(w, h, 51), synth.flat_image(w, h)])
    L = orbfe.Extractor(*params, w, h, max_batch=5)
    R = orbfe.Extractor(*params, w, h, max_batch=5)
    kl, dl, nl = L.extract_batch(lefts)
    kr, dr, nr = R.extract_batch(rights)
    mb = np.float32(s["bf"]) / np.float32(s["fx"])
    ur, dp, nm = orbfe.stereo_match(L, R, 5, np.float32(s["bf"]), mb)
    oL, oR = oracle.Extractor(*params), oracle.Extractor(*params)
    for b in range(5):
        okl, odl = oL.extract(lefts[b])
        okr, odr = oR.extract(rights[b])
        our, odp, on = oracle.stereo_matches(oL, oR, okl, odl, okr, odr, np.float32(s["bf"]), mb)
        assert nl[b] == len(okl) and nr[b] == len(okr)
        assert np.array_equal(ur[b][:nl[b]], our), "image %d" % b
        assert np.array_equal(dp[b][:nl[b]], odp), "image %d" % b
        assert nm[b] == on


def test_stereo_frame_one_call(oracle):
    """orbfe_stereo_frame (the stereo Frame constructor, src/Frame.cc:61-116, in one C-ABI call) against the oracle,
    batch of 2, called twice on the same handles (the second call reuses the captured graphs)."""
    from my_orb_slam2_b200 import orbfe
    w, h = synth.SHAPES["kitti"]
    s = synth.SETTINGS["kitti"]
    params = (s["nfeatures"], s["scale_factor"], s["nlevels"], s["ini_th"], s["min_th"])
    pairs = [synth.synth_pair(w, h, 70 + i) for i in range(2)]
    lefts, rights = np.stack([p[0] for p in pairs]), np.stack([p[1] for p in pairs])
    mbf, mb = np.float32(s["bf"]), np.float32(s["bf"]) / np.float32(s["fx"])
    oL, oR = oracle.Extractor(*params), oracle.Extractor(*params)
    want = []
    for b in range(2):
        okl, odl = oL.extract(lefts[b])
        okr, odr = oR.extract(rights[b])
        want.append((okl, odl, okr, odr) + tuple(oracle.stereo_matches(oL, oR, okl, odl, okr, odr, mbf, mb)))
    # max_batch == batch: results leave as one block per handle and are scattered on the host; max_batch > batch: the
    # general path (separate copies straight into the caller's arrays)
    for max_batch in (2, 2, 3):
        if max_batch == 3 or "L" not in locals():
            L = orbfe.Extractor(*params, w, h, max_batch=max_batch)
            R = orbfe.Extractor(*params, w, h, max_batch=max_batch)
        kl, dl, nl, kr, dr, nr, ur, dp, nm = orbfe.stereo_frame(L, R, lefts, rights, mbf, mb)
        for b in range(2):
            okl, odl, okr, odr, our, odp, on = want[b]
            assert nl[b] == len(okl) and nr[b] == len(okr)
            assert np.array_equal(kl[b][:nl[b]], okl) and np.array_equal(dl[b][:nl[b]], odl)
            assert np.array_equal(kr[b][:nr[b]], okr) and np.array_equal(dr[b][:nr[b]], odr)
            assert np.array_equal(ur[b][:nl[b]], our) and np.array_equal(dp[b][:nl[b]], odp)
            assert nm[b] == on and on > 50
