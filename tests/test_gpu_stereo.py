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


def test_stereo_pipelined_pairs_without_host_sync():
    """Two DIFFERENT batches of pairs queued back to back on the two handles' streams with no host synchronisation in
    between (the bench's device-resident loop): the matcher of step k reads the right handle's pyramid / keypoints /
    descriptors on the LEFT stream, so the right handle's upload + run of step k+1 must wait for it
    (orbfe_stereo_match_run orders the right stream after the stereo kernels).  Results must equal the synchronous path."""
    import torch
    from my_orb_slam2_b200 import orbfe
    w, h = synth.SHAPES["kitti"]
    s = synth.SETTINGS["kitti"]
    params = (s["nfeatures"], s["scale_factor"], s["nlevels"], s["ini_th"], s["min_th"])
    mbf, mb = np.float32(s["bf"]), np.float32(s["bf"]) / np.float32(s["fx"])
    B, steps = 8, 4
    dev = torch.device("cuda", 0)
    sets = []
    for k in range(steps):
        pairs = [synth.synth_pair(w, h, 300 + 10 * k + i) for i in range(2)]
        idx = np.arange(B) % 2
        sets.append((torch.from_numpy(np.stack([p[0] for p in pairs])[idx]).to(dev),
                     torch.from_numpy(np.stack([p[1] for p in pairs])[idx]).to(dev)))
    L = orbfe.Extractor(*params, w, h, max_batch=B)
    R = orbfe.Extractor(*params, w, h, max_batch=B)
    cap = L.capacity
    lib = orbfe.lib()

    def run(sync_every_step):
        outs = []
        for dl, dr in sets:
            ur = torch.empty((B, cap), dtype=torch.float32).pin_memory()
            dp = torch.empty((B, cap), dtype=torch.float32).pin_memory()
            nm = torch.zeros(B, dtype=torch.int32).pin_memory()
            L.upload(dl.data_ptr(), B, on_device=True)
            R.upload(dr.data_ptr(), B, on_device=True)
            L.run(B)
            R.run(B)
            orbfe.check(lib.orbfe_stereo_match_run(L._h, R._h, B, mbf, mb))
            orbfe.check(lib.orbfe_stereo_match_download_async(L._h, B, ur.data_ptr(), dp.data_ptr(), cap, nm.data_ptr()))
            if sync_every_step:
                L.sync()
                R.sync()
            outs.append((ur, dp, nm))
        L.sync()
        R.sync()
        return [(a.numpy().copy(), b.numpy().copy(), c.numpy().copy()) for a, b, c in outs]

    want = run(True)
    for _ in range(3):
        got = run(False)
        for k in range(steps):
            assert np.array_equal(got[k][2], want[k][2]), "step %d match counts" % k
            # entries past an image's keypoint count are -1 (the reference initialises all N to -1)
            assert np.array_equal(got[k][0], want[k][0]) and np.array_equal(got[k][1], want[k][1]), "step %d" % k
    assert want[0][2].min() > 50 and not np.array_equal(want[0][0], want[1][0])
