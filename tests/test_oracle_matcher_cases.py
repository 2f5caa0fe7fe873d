"""The synthetic matcher / stereo workloads are meaningful for the oracle:
matches exist, claims collide, the rotation histogram and the median cut remove something."""
import numpy as np

from cases import extract_with_oracle, views
from my_orb_slam2_b200 import synth


def test_stereo_case_has_matches(oracle):
    params, (left, right), (exl, exr), ((kl, dl), (kr, dr)) = extract_with_oracle(oracle, "kitti", 0, pair=True)
    s = synth.SETTINGS["kitti"]
    ur, dp, n = oracle.stereo_matches(exl, exr, kl, dl, kr, dr, s["bf"], s["bf"] / s["fx"])
    assert 1900 <= len(kl) <= 2100
    assert n > 150 and (ur >= 0).sum() == n and np.all(dp[ur >= 0] > 0)


def test_projection_cases(oracle):
    params, img, ex, (kp, desc) = extract_with_oracle(oracle, "kitti", 0)
    frame = synth.frame_arrays(kp, desc, (0.0, 0.0, 1241.0, 376.0))
    mps = synth.map_points_case(frame, 20000, 1500, seed=0)
    fv, mv = views(oracle, frame=frame, mps=mps)
    assigned, n = oracle.search_by_projection_map(fv, mv, 1.0, 0.8)
    assert n > 500 and (assigned >= 0).sum() <= n
    last, cam = synth.last_frame_case(frame, seed=0)
    fv, lv, cv = views(oracle, frame=frame, last=last, cam=cam)
    a1, n1 = oracle.search_by_projection_last(fv, cv, lv, 7.0, False, True)
    a0, n0 = oracle.search_by_projection_last(fv, cv, lv, 7.0, False, False)
    assert n0 > 500 and n1 < n0  # the rotation histogram removed outliers


def test_triangulation_case(oracle):
    params, img, ex, (kp, desc) = extract_with_oracle(oracle, "kitti", 0)
    k1, k2, F12, ep = synth.keyframe_pair_case(kp, desc, seed=0)
    v1 = views(oracle, kf=k1)
    v2 = views(oracle, kf=k2)
    pairs = oracle.search_for_triangulation(v1, v2, F12, float(ep[0]), float(ep[1]), False, False)
    pairs_ori = oracle.search_for_triangulation(v1, v2, F12, float(ep[0]), float(ep[1]), False, True)
    assert len(pairs) > 300 and len(pairs_ori) <= len(pairs)
    assert np.all(np.diff(pairs[:, 0]) > 0)
