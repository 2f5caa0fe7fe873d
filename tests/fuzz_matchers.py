"""Randomised parity cases for the ORBmatcher members: CUDA path (C ABI through my_orb_slam2_b200.frontend) against
the CPU oracle.  Not collected by pytest itself: tests/test_gpu_fuzz.py runs a bounded number of seeds, and
`python tests/fuzz_matchers.py --first 0 --count 200` runs a long sweep on a GPU box.

Per seed: a random subset of the features of one synthetic image, then every matcher member once with drawn
thresholds, ratios, window sizes, motion and scale (src/ORBmatcher.cc: SearchByProjection x4, SearchByBoW x2,
SearchForInitialization, SearchForTriangulation, Fuse x2, SearchBySim3; Frame::isInFrustum + SearchByProjection).
All outputs are indices / counts / copied coordinates: compared bit for bit."""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

BOUNDS = (0.0, 0.0, 1241.0, 376.0)
_features = {}


def features(oracle, seed):
    from cases import extract_with_oracle
    if seed not in _features:
        _features[seed] = extract_with_oracle(oracle, "kitti", seed)[3]
    return _features[seed]


def _same(a, b):
    a, b = np.asarray(a), np.asarray(b)
    if a.dtype == np.float32:
        a, b = a.view(np.uint32), b.view(np.uint32)
    return a.shape == b.shape and np.array_equal(a, b)


def run_case(seed, oracle, fe):
    """Returns a list of (member, status, matches the oracle found) with status "ok" or "DIFF"."""
    from cases import points_view, projection, sim3_pair, views
    from my_orb_slam2_b200 import synth
    rng = np.random.default_rng([seed, 4242])
    kp, desc = features(oracle, int(rng.integers(0, 3)))
    n = int(rng.choice([40, 200, 700, len(kp), len(kp)]))
    sel = np.sort(rng.choice(len(kp), min(n, len(kp)), replace=False))
    kp, desc = kp[sel], desc[sel]
    out = []

    def put(name, ok, n_found):
        out.append((name, "ok" if ok else "DIFF", int(n_found)))

    ratio = float(rng.choice([0.6, 0.7, 0.75, 0.8, 0.9, 0.95]))
    ori = bool(rng.integers(0, 2))
    # SearchByProjection(Frame, MapPoints)  (:46-140)
    frame = synth.frame_arrays(kp, desc, BOUNDS, seed=seed, taken_frac=float(rng.choice([0.0, 0.05, 0.3])))
    n_map = int(rng.choice([50, 1000, 5000, 20000]))
    mps = synth.map_points_case(frame, n_map, min(n_map, int(rng.choice([20, 500, 1500, 5000]))), seed=seed)
    th = float(rng.choice([1.0, 3.0, 5.0, 9.0]))
    want, wn = oracle.search_by_projection_map(*views(oracle, frame=frame, mps=mps), th, ratio)
    got, gn = fe.ORBmatcher(ratio).SearchByProjection(*views(fe, frame=frame, mps=mps), th)
    put("SearchByProjection(map)", gn == wn and _same(got, want), wn)
    # SearchByProjection(Cur, Last)  (:1392-1538)
    tz, mono = float(rng.choice([0.0, 0.3, 2.0, -2.0])), bool(rng.integers(0, 2))
    th = float(rng.choice([7.0, 15.0, 30.0]))
    last, cam = synth.last_frame_case(frame, seed=seed, tz=tz)
    ofv, olv, ocv = views(oracle, frame=frame, last=last, cam=cam)
    want, wn = oracle.search_by_projection_last(ofv, ocv, olv, th, mono, ori)
    fv, lv, cv = views(fe, frame=frame, last=last, cam=cam)
    got, gn = fe.ORBmatcher(ratio, ori).SearchByProjection(fv, lv, th, bMono=mono, cam=cv)
    put("SearchByProjection(last)", gn == wn and _same(got, want), wn)
    # SearchForTriangulation  (:702-872)
    k1, k2, F12, ep = synth.keyframe_pair_case(kp, desc, seed=seed, n_nodes=int(rng.choice([1, 10, 100, 1000])))
    only_stereo = bool(rng.integers(0, 2))
    want = oracle.search_for_triangulation(views(oracle, kf=k1), views(oracle, kf=k2), F12, float(ep[0]), float(ep[1]),
                                           only_stereo, ori)
    got = fe.ORBmatcher(ratio, ori).SearchForTriangulation(views(fe, kf=k1), views(fe, kf=k2), F12, ep, only_stereo)
    put("SearchForTriangulation", _same(got, want), len(want))
    # SearchByBoW x2  (:182-319, :563-696)
    k1, k2 = synth.bow_case(kp, desc, seed=seed, n_nodes=int(rng.choice([1, 7, 100, 1500])))
    for kf_kf in (False, True):
        want, wn = oracle.search_by_bow(views(oracle, kf=k1), views(oracle, kf=k2), kf_kf, ratio, ori)
        got, gn = fe.ORBmatcher(ratio, ori).SearchByBoW(views(fe, kf=k1), views(fe, kf=k2), kf_kf)
        put("SearchByBoW(%s)" % ("KF,KF" if kf_kf else "KF,F"), gn == wn and _same(got, want), wn)
    # SearchForInitialization  (:446-561)
    window = int(rng.choice([10, 30, 100, 400]))
    f1, f2, prev = synth.initialization_case(kp, desc, BOUNDS, seed=seed, shift=float(rng.choice([0.0, 5.0, 12.0, 40.0])))
    wm, wp, wn = oracle.search_for_initialization(views(oracle, frame=f1), views(oracle, frame=f2), prev, window, ratio,
                                                  ori)
    gm, gp, gn = fe.ORBmatcher(ratio, ori).SearchForInitialization(views(fe, frame=f1), views(fe, frame=f2), prev,
                                                                   window)
    put("SearchForInitialization", gn == wn and _same(gm, wm) and _same(gp, wp), wn)
    # SearchByProjection reloc / Sim3  (:1540-1667, :321-434), Fuse x2 (:879-1156)
    frame = synth.frame_arrays(kp, desc, BOUNDS, seed=seed + 1, taken_frac=float(rng.choice([0.0, 0.1, 0.4])))
    n_extra = int(rng.choice([0, 400, 5000]))
    pts, pr = synth.points3d_case(frame, seed=seed, n_extra=n_extra, noise_px=float(rng.choice([0.5, 2.0, 6.0])))
    th, od = float(rng.choice([3.0, 10.0, 25.0])), int(rng.choice([50, 64, 100]))
    want, wn = oracle.search_by_projection_reloc(views(oracle, frame=frame), projection(oracle, pr),
                                                 points_view(oracle, pts), th, od, ori)
    got, gn = fe.ORBmatcher(ratio, ori).SearchByProjectionReloc(views(fe, frame=frame), projection(fe, pr),
                                                               points_view(fe, pts), th, od)
    put("SearchByProjection(reloc)", gn == wn and _same(got, want), wn)
    sig = (1.0 / synth.scale_factors(8) ** 2).astype(np.float32)
    ow = (-pr["tcw"][:, :3].T @ pr["tcw"][:, 3]).astype(np.float32)
    th = float(rng.choice([3.0, 4.0, 12.0]))
    want, wn = oracle.fuse(views(oracle, frame=frame), projection(oracle, pr, ow=ow), points_view(oracle, pts), th, sig)
    got, gn = fe.ORBmatcher(0.6, True).Fuse(views(fe, frame=frame), projection(fe, pr, ow=ow), points_view(fe, pts), th,
                                           sig)
    put("Fuse(KF,MPs)", gn == wn and _same(got, want), wn)
    obs = rng.integers(0, 9, len(pts["valid"])).astype(np.int32)
    th = float(rng.choice([1.0, 3.0, 5.0]))
    wa, wn, wf = oracle.search_local_points(views(oracle, frame=frame), projection(oracle, pr), points_view(oracle, pts),
                                            obs, 0.5, th, ratio)
    ga, gn, gf = fe.ORBmatcher(ratio, True).SearchLocalPoints(views(fe, frame=frame), projection(fe, pr),
                                                             points_view(fe, pts), obs, th)
    put("isInFrustum+SearchByProjection", gn == wn and _same(ga, wa) and all(_same(gf[k], wf[k]) for k in wf), wn)
    scale = float(rng.choice([0.8, 0.95, 1.0, 1.07, 1.6]))
    pts, pr = synth.points3d_case(frame, seed=seed, sim3_scale=scale, n_extra=n_extra)
    th = float(rng.choice([4, 10, 30]))
    want, wn = oracle.search_by_projection_sim3(views(oracle, frame=frame), projection(oracle, pr),
                                                points_view(oracle, pts), th)
    got, gn = fe.ORBmatcher(0.75, True).SearchByProjectionSim3(views(fe, frame=frame), projection(fe, pr),
                                                              points_view(fe, pts), th)
    put("SearchByProjection(Sim3)", gn == wn and _same(got, want), wn)
    want, wn = oracle.fuse(views(oracle, frame=frame), projection(oracle, pr), points_view(oracle, pts), th, None)
    got, gn = fe.ORBmatcher(0.6, True).Fuse(views(fe, frame=frame), projection(fe, pr), points_view(fe, pts), th, None)
    put("Fuse(KF,Scw,MPs)", gn == wn and _same(got, want), wn)
    # SearchBySim3  (:1158-1390)
    f1, f2, p1, p2, pair = synth.sim3_pair_case(frame, seed=seed, s12=float(rng.choice([0.9, 1.0, 1.03, 1.2])))
    th = float(rng.choice([3.0, 7.5, 20.0]))
    want, wn = oracle.search_by_sim3(views(oracle, frame=f1), views(oracle, frame=f2), points_view(oracle, p1),
                                     points_view(oracle, p2), sim3_pair(oracle, pair), th)
    got, gn = fe.ORBmatcher(0.75, True).SearchBySim3(views(fe, frame=f1), views(fe, frame=f2), points_view(fe, p1),
                                                    points_view(fe, p2), sim3_pair(fe, pair), th)
    put("SearchBySim3", gn == wn and _same(got, want), wn)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--first", type=int, default=0)
    ap.add_argument("--count", type=int, default=100)
    args = ap.parse_args()
    from oracle import pyoracle as oracle
    from my_orb_slam2_b200 import frontend
    oracle.build()
    tally, bad = {}, 0
    for seed in range(args.first, args.first + args.count):
        for name, st, found in run_case(seed, oracle, frontend):
            t = tally.setdefault(name, [0, 0, 0, 0])
            t[0 if st == "ok" else 1] += 1
            t[2] += found
            t[3] += found == 0
            if st != "ok":
                bad += 1
                print("FAIL seed %d %s" % (seed, name), flush=True)
    for name, (ok, diff, found, empty) in tally.items():
        print("%-34s %5d equal %5d different | %9d matches in all, %4d cases without any" % (name, ok, diff, found, empty))
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
