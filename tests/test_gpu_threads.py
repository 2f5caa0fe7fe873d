"""Re-entrancy of the C ABI (SURVEY.md 8(b) "Threading"): the reference drives the left / right extractors from two
std::threads per stereo frame (src/Frame.cc:89-92) and uses stack-local ORBmatcher objects concurrently from the
Tracking, LocalMapping and LoopClosing threads.  Here host threads (ctypes drops the GIL inside every call) hammer
their own extractor / matcher handles at the same time; every result must equal the one computed alone."""
import threading

import numpy as np
import pytest

from cases import extract_with_oracle, points_view, projection, views
from my_orb_slam2_b200 import synth

pytestmark = pytest.mark.gpu
BOUNDS = (0.0, 0.0, 1241.0, 376.0)


def _run_threads(workers):
    errors = []

    def guard(fn):
        def run():
            try:
                fn()
            except BaseException as e:  # noqa: BLE001 - reported below, in the main thread
                errors.append(repr(e))
        return run

    ts = [threading.Thread(target=guard(w)) for w in workers]
    for t in ts:
        t.start()
    for t in ts:
        t.join(300)
    assert not any(t.is_alive() for t in ts), "a worker thread hangs"
    assert not errors, errors


def test_extractors_and_matchers_from_concurrent_threads(oracle):
    from my_orb_slam2_b200 import frontend as fe
    shapes = [("kitti", 1241, 376), ("euroc", 752, 480), ("tum", 640, 480), ("kitti", 1241, 376)]
    images, params, alone = [], [], []
    for i, (name, w, h) in enumerate(shapes):
        s = synth.SETTINGS[name]
        params.append((s["nfeatures"], s["scale_factor"], s["nlevels"], s["ini_th"], s["min_th"]))
        images.append([synth.synth_image(w, h, 100 + 10 * i + j) for j in range(3)])
        ex = fe.ORBextractor(*params[-1])
        alone.append([ex(img) for img in images[-1]])
    # matcher work: three different members, each on its own matcher object
    kp, desc = extract_with_oracle(oracle, "kitti", 0)[3]
    frame = synth.frame_arrays(kp, desc, BOUNDS, seed=5, taken_frac=0.1)
    mps = synth.map_points_case(frame, 20000, 1500, seed=5)
    pts, pr = synth.points3d_case(frame, seed=5)
    k1, k2 = synth.bow_case(kp, desc, seed=5)

    def m_proj():
        return fe.ORBmatcher(0.8).SearchByProjection(*views(fe, frame=frame, mps=mps), 3.0)

    def m_bow():
        return fe.ORBmatcher(0.7, True).SearchByBoW(views(fe, kf=k1), views(fe, kf=k2), False)

    def m_reloc():
        return fe.ORBmatcher(0.9, True).SearchByProjectionReloc(views(fe, frame=frame), projection(fe, pr),
                                                               points_view(fe, pts), 10.0, 100)

    members = [m_proj, m_bow, m_reloc]
    m_alone = [m() for m in members]
    assert all(n > 100 for _, n in m_alone)
    rounds = 12

    def extractor_worker(i):
        def run():
            ex = fe.ORBextractor(*params[i])  # created inside the thread, like Tracking's constructor would
            for r in range(rounds):
                j = r % 3
                k, d = ex(images[i][j])
                assert np.array_equal(k, alone[i][j][0]) and np.array_equal(d, alone[i][j][1]), (i, r)
        return run

    def matcher_worker(i):
        def run():
            for r in range(rounds * 2):
                got, n = members[i]()
                assert n == m_alone[i][1] and np.array_equal(got, m_alone[i][0]), (i, r)
        return run

    _run_threads([extractor_worker(i) for i in range(4)] + [matcher_worker(i) for i in range(3)])


def test_stereo_pair_extracted_on_two_threads_then_matched(oracle):
    """The stereo Frame constructor's schedule: ExtractORB(0) and ExtractORB(1) on two threads, join, then
    ComputeStereoMatches on the calling thread (src/Frame.cc:89-102)."""
    from my_orb_slam2_b200 import frontend as fe, orbfe
    s = synth.SETTINGS["kitti"]
    w, h = synth.SHAPES["kitti"]
    params = (s["nfeatures"], s["scale_factor"], s["nlevels"], s["ini_th"], s["min_th"])
    eL, eR = fe.ORBextractor(*params), fe.ORBextractor(*params)
    oL, oR = oracle.Extractor(*params), oracle.Extractor(*params)
    mb = np.float32(s["bf"]) / np.float32(s["fx"])
    for seed in range(4):
        left, right = synth.synth_pair(w, h, 200 + seed)
        res = {}
        _run_threads([lambda: res.__setitem__("l", eL(left)), lambda: res.__setitem__("r", eR(right))])
        ur, dp, n = orbfe.stereo_match(eL._h, eR._h, 1, np.float32(s["bf"]), mb)
        kl, dl = oL.extract(left)
        kr, dr = oR.extract(right)
        our, odp, on = oracle.stereo_matches(oL, oR, kl, dl, kr, dr, np.float32(s["bf"]), mb)
        assert np.array_equal(res["l"][0], kl) and np.array_equal(res["l"][1], dl)
        assert np.array_equal(res["r"][0], kr) and np.array_equal(res["r"][1], dr)
        assert np.array_equal(ur[0][:len(kl)], our) and np.array_equal(dp[0][:len(kl)], odp) and n[0] == on


def test_one_matcher_handle_from_two_threads_fails_loudly(oracle):
    """Two threads inside ONE orbfe_matcher: a call either completes with the right answer or is refused with
    ORBFE_ERR_STATE — never a silently wrong result (the handle's staging arena would be overwritten)."""
    from my_orb_slam2_b200 import orbfe
    m = orbfe.Matcher(0)
    q, mp = synth.bruteforce_case(nq=2048, nm=60000, n_planted=500, seed=9)
    want_b, want_s = m.hamming_bruteforce(q, mp)
    stats = {"ok": 0, "refused": 0}
    lock = threading.Lock()

    def worker():
        for _ in range(30):
            try:
                b, s = m.hamming_bruteforce(q, mp)
            except orbfe.OrbfeError as e:
                assert e.code == orbfe.ERR_STATE, e
                with lock:
                    stats["refused"] += 1
                continue
            assert np.array_equal(b, want_b) and np.array_equal(s, want_s)
            with lock:
                stats["ok"] += 1

    _run_threads([worker, worker, worker])
    assert stats["ok"] >= 1 and stats["ok"] + stats["refused"] == 90
    print("one handle, three threads: %(ok)d calls completed, %(refused)d refused" % stats)
