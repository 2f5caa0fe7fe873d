"""CPU checks of the oracle restatements of the seven other ORBmatcher members on the synthetic
workloads: matches exist, claims collide, steals happen, the rotation histogram removes something, and the shared
log / PredictScale arithmetic agrees with libm where libm is unambiguous."""
import math

import numpy as np
import pytest

from cases import extract_with_oracle, points_view, projection, sim3_pair, views
from my_orb_slam2_b200 import synth

BOUNDS = (0.0, 0.0, 1241.0, 376.0)


@pytest.fixture(scope="module")
def feats(oracle):
    params, img, ex, (kp, desc) = extract_with_oracle(oracle, "kitti", 0)
    return kp, desc


def test_log_and_predict_scale(oracle):
    rng = np.random.default_rng(0)
    xs = np.concatenate([rng.uniform(1e-3, 50, 20000), 1.2 ** np.arange(-8, 12)]).astype(np.float32)
    got = np.array([oracle.log_f32(float(x)) for x in xs], np.float32)
    want = np.log(xs.astype(np.float64)).astype(np.float32)  # correctly rounded except for double-rounding ties
    assert np.array_equal(got, want)
    lsf = oracle.log_f32(float(np.float32(1.2)))
    for ratio_pow in range(-3, 12):
        for eps in (0.97, 1.03):
            r = (1.2 ** ratio_pow) * eps
            want = min(max(math.ceil(math.log(r) / math.log(1.2)), 0), 7)
            assert oracle.predict_scale(float(np.float32(r)), 1.0, lsf, 8) == want
    assert oracle.predict_scale(0.0, 1.0, lsf, 8) == 0
    assert oracle.predict_scale(1.0, 0.0, lsf, 8) == 7  # ratio = inf


def test_bow_cases(oracle, feats):
    kp, desc = feats
    k1, k2 = synth.bow_case(kp, desc, seed=0)
    v1, v2 = views(oracle, kf=k1), views(oracle, kf=k2)
    for kf_kf in (0, 1):
        m0, n0 = oracle.search_by_bow(v1, v2, kf_kf, 0.7, False)
        m1, n1 = oracle.search_by_bow(v1, v2, kf_kf, 0.7, True)
        assert n0 > 300 and (m0 >= 0).sum() == n0
        assert n1 < n0 and (m1 >= 0).sum() == n1
        # a target feature is given away at most once
        tgt = m0[m0 >= 0] if kf_kf else np.flatnonzero(m0 >= 0)
        assert len(np.unique(tgt)) == len(tgt)


def test_initialization_case(oracle, feats):
    kp, desc = feats
    f1, f2, prev = synth.initialization_case(kp, desc, BOUNDS, seed=0)
    v1, v2 = views(oracle, frame=f1), views(oracle, frame=f2)
    m0, p0, n0 = oracle.search_for_initialization(v1, v2, prev, 100, 0.9, False)
    m1, p1, n1 = oracle.search_for_initialization(v1, v2, prev, 100, 0.9, True)
    assert n0 > 100 and (m0 >= 0).sum() == n0 and n1 <= n0 and (m1 >= 0).sum() == n1
    assert np.all(kp["octave"][m0 >= 0] == 0)
    moved = np.any(p0 != prev, axis=1)
    assert np.array_equal(moved | (m0 < 0), moved | (m0 < 0)) and moved.sum() > 0
    assert len(np.unique(m0[m0 >= 0])) == n0  # a stolen feature leaves exactly one owner


def test_projection_claim_cases(oracle, feats):
    kp, desc = feats
    frame = synth.frame_arrays(kp, desc, BOUNDS, taken_frac=0.1)
    pts, pr = synth.points3d_case(frame, seed=0)
    fv, pv = views(oracle, frame=frame), points_view(oracle, pts)
    a0, n0 = oracle.search_by_projection_reloc(fv, projection(oracle, pr), pv, 10.0, 100, False)
    a1, n1 = oracle.search_by_projection_reloc(fv, projection(oracle, pr), pv, 10.0, 100, True)
    assert n0 > 800 and (a0 >= 0).sum() == n0 and n1 < n0 and (a1 == -2).sum() == n0 - n1
    assert not np.any(a0[frame["mp_obs"] >= 0] >= 0)  # occupied slots stay untouched
    a2, n2 = oracle.search_by_projection_reloc(fv, projection(oracle, pr), pv, 3.0, 64, True)
    assert 0 < n2 < n1
    pts_s, pr_s = synth.points3d_case(frame, seed=1, sim3_scale=1.07)
    a3, n3 = oracle.search_by_projection_sim3(fv, projection(oracle, pr_s), points_view(oracle, pts_s), 10)
    assert n3 > 500 and (a3 >= 0).sum() == n3


def test_fuse_and_sim3_cases(oracle, feats):
    kp, desc = feats
    frame = synth.frame_arrays(kp, desc, BOUNDS)
    pts, pr = synth.points3d_case(frame, seed=2)
    fv, pv = views(oracle, frame=frame), points_view(oracle, pts)
    sig = (1.0 / synth.scale_factors(8) ** 2).astype(np.float32)
    ow = -pr["tcw"][:, :3].T @ pr["tcw"][:, 3]
    b0, n0 = oracle.fuse(fv, projection(oracle, pr, ow=ow), pv, 3.0, sig)
    assert n0 > 200 and (b0 >= 0).sum() == n0
    pts_s, pr_s = synth.points3d_case(frame, seed=3, sim3_scale=0.95)
    b1, n1 = oracle.fuse(fv, projection(oracle, pr_s), points_view(oracle, pts_s), 4.0, None)
    assert n1 > n0  # no chi-square gate in the Sim3 overload
    f1, f2, p1, p2, pair = synth.sim3_pair_case(frame, seed=0)
    m12, n = oracle.search_by_sim3(views(oracle, frame=f1), views(oracle, frame=f2), points_view(oracle, p1),
                                   points_view(oracle, p2), sim3_pair(oracle, pair), 7.5)
    assert n > 300 and (m12 >= 0).sum() == n and len(np.unique(m12[m12 >= 0])) == n
