"""GPU parity of the seven other ORBmatcher members against the CPU oracle: SearchByBoW x2,
SearchForInitialization, the relocalisation and Sim3 SearchByProjection overloads, Fuse x2, SearchBySim3.
All outputs are indices / counts (and, for SearchForInitialization, copied keypoint coordinates): bit-exact."""
import numpy as np
import pytest

from cases import extract_with_oracle, points_view, projection, sim3_pair, views
from my_orb_slam2_b200 import synth

pytestmark = pytest.mark.gpu
BOUNDS = (0.0, 0.0, 1241.0, 376.0)


@pytest.fixture(scope="module")
def fe():
    from my_orb_slam2_b200 import frontend
    return frontend


@pytest.fixture(scope="module")
def feats(oracle):
    params, img, ex, (kp, desc) = extract_with_oracle(oracle, "kitti", 0)
    return kp, desc


@pytest.mark.parametrize("kf_kf", [False, True])
@pytest.mark.parametrize("seed,n_nodes,ratio,ori", [(0, 100, 0.7, True), (1, 100, 0.75, False), (2, 7, 0.9, True),
                                                    (3, 1, 0.9, True), (4, 1500, 0.6, True)])
def test_search_by_bow(fe, oracle, feats, kf_kf, seed, n_nodes, ratio, ori):
    kp, desc = feats
    k1, k2 = synth.bow_case(kp, desc, seed=seed, n_nodes=n_nodes)
    want, wn = oracle.search_by_bow(views(oracle, kf=k1), views(oracle, kf=k2), kf_kf, ratio, ori)
    got, gn = fe.ORBmatcher(ratio, ori).SearchByBoW(views(fe, kf=k1), views(fe, kf=k2), kf_kf)
    assert gn == wn and np.array_equal(got, want)
    assert wn > 50


def test_search_by_bow_empty(fe, feats):
    kp, desc = feats
    k1, k2 = synth.bow_case(kp, desc, seed=0)
    k1 = dict(k1, has_mp=np.zeros_like(k1["has_mp"]))
    got, gn = fe.ORBmatcher(0.7, True).SearchByBoW(views(fe, kf=k1), views(fe, kf=k2), False)
    assert gn == 0 and np.all(got == -1)


@pytest.mark.parametrize("seed,window,ratio,ori", [(0, 100, 0.9, True), (1, 100, 0.9, False), (2, 30, 0.8, True),
                                                   (3, 400, 0.95, True)])
def test_search_for_initialization(fe, oracle, feats, seed, window, ratio, ori):
    kp, desc = feats
    f1, f2, prev = synth.initialization_case(kp, desc, BOUNDS, seed=seed)
    wm, wp, wn = oracle.search_for_initialization(views(oracle, frame=f1), views(oracle, frame=f2), prev, window,
                                                  ratio, ori)
    gm, gp, gn = fe.ORBmatcher(ratio, ori).SearchForInitialization(views(fe, frame=f1), views(fe, frame=f2), prev,
                                                                   window)
    assert gn == wn and np.array_equal(gm, wm)
    assert np.array_equal(gp.view(np.uint32), wp.view(np.uint32))
    assert wn > 20


@pytest.mark.parametrize("seed,th,orb_dist,ori", [(0, 10.0, 100, True), (1, 3.0, 64, True), (2, 10.0, 100, False),
                                                  (3, 25.0, 100, True)])
def test_search_by_projection_reloc(fe, oracle, feats, seed, th, orb_dist, ori):
    kp, desc = feats
    frame = synth.frame_arrays(kp, desc, BOUNDS, seed=seed, taken_frac=0.1)
    pts, pr = synth.points3d_case(frame, seed=seed)
    want, wn = oracle.search_by_projection_reloc(views(oracle, frame=frame), projection(oracle, pr),
                                                 points_view(oracle, pts), th, orb_dist, ori)
    got, gn = fe.ORBmatcher(0.9, ori).SearchByProjectionReloc(views(fe, frame=frame), projection(fe, pr),
                                                             points_view(fe, pts), th, orb_dist)
    assert gn == wn and np.array_equal(got, want)
    assert wn > 100


@pytest.mark.parametrize("seed,th,scale", [(0, 10, 1.07), (1, 10, 0.9), (2, 4, 1.0), (3, 30, 2.5)])
def test_search_by_projection_sim3(fe, oracle, feats, seed, th, scale):
    kp, desc = feats
    frame = synth.frame_arrays(kp, desc, BOUNDS, seed=seed, taken_frac=0.2)
    pts, pr = synth.points3d_case(frame, seed=seed, sim3_scale=scale)
    want, wn = oracle.search_by_projection_sim3(views(oracle, frame=frame), projection(oracle, pr),
                                                points_view(oracle, pts), th)
    got, gn = fe.ORBmatcher(0.75, True).SearchByProjectionSim3(views(fe, frame=frame), projection(fe, pr),
                                                              points_view(fe, pts), th)
    assert gn == wn and np.array_equal(got, want)
    assert wn > 100


@pytest.mark.parametrize("seed,th,scale", [(0, 3.0, None), (1, 3.0, None), (2, 4.0, 0.95), (3, 4.0, 1.3),
                                           (4, 12.0, None)])
def test_fuse(fe, oracle, feats, seed, th, scale):
    kp, desc = feats
    frame = synth.frame_arrays(kp, desc, BOUNDS, seed=seed)
    pts, pr = synth.points3d_case(frame, seed=seed, sim3_scale=scale)
    sig = (1.0 / synth.scale_factors(8) ** 2).astype(np.float32) if scale is None else None
    ow = None if scale is not None else (-pr["tcw"][:, :3].T @ pr["tcw"][:, 3]).astype(np.float32)
    want, wn = oracle.fuse(views(oracle, frame=frame), projection(oracle, pr, ow=ow), points_view(oracle, pts), th,
                           sig)
    got, gn = fe.ORBmatcher(0.6, True).Fuse(views(fe, frame=frame), projection(fe, pr, ow=ow), points_view(fe, pts),
                                           th, sig)
    assert gn == wn and np.array_equal(got, want)
    assert wn > 100


@pytest.mark.parametrize("seed,th,s12", [(0, 7.5, 1.03), (1, 7.5, 0.97), (2, 3.0, 1.0), (3, 20.0, 1.2)])
def test_search_by_sim3(fe, oracle, feats, seed, th, s12):
    kp, desc = feats
    frame = synth.frame_arrays(kp, desc, BOUNDS, seed=seed)
    f1, f2, p1, p2, pair = synth.sim3_pair_case(frame, seed=seed, s12=s12)
    want, wn = oracle.search_by_sim3(views(oracle, frame=f1), views(oracle, frame=f2), points_view(oracle, p1),
                                     points_view(oracle, p2), sim3_pair(oracle, pair), th)
    got, gn = fe.ORBmatcher(0.75, True).SearchBySim3(views(fe, frame=f1), views(fe, frame=f2), points_view(fe, p1),
                                                    points_view(fe, p2), sim3_pair(fe, pair), th)
    assert gn == wn and np.array_equal(got, want)
    assert wn > 50


def test_empty_inputs(fe, feats):
    kp, desc = feats
    frame = synth.frame_arrays(kp, desc, BOUNDS)
    pts, pr = synth.points3d_case(frame, seed=0)
    none = {k: (v[:0] if v is not None else None) for k, v in pts.items()}
    got, gn = fe.ORBmatcher(0.9, True).SearchByProjectionReloc(views(fe, frame=frame), projection(fe, pr),
                                                              points_view(fe, none), 10.0, 100)
    assert gn == 0 and np.all(got == -1)
    got, gn = fe.ORBmatcher(0.6, True).Fuse(views(fe, frame=frame), projection(fe, pr), points_view(fe, none), 3.0,
                                           np.ones(8, np.float32))
    assert gn == 0 and len(got) == 0


@pytest.mark.parametrize("seed,th,n_extra", [(0, 1.0, 400), (1, 3.0, 400), (2, 5.0, 18000), (3, 1.0, 0)])
def test_search_local_points(fe, oracle, feats, seed, th, n_extra):
    """Frame::isInFrustum over the local map fused with SearchByProjection."""
    kp, desc = feats
    frame = synth.frame_arrays(kp, desc, BOUNDS, seed=seed, taken_frac=0.1)
    pts, pr = synth.points3d_case(frame, seed=seed, n_extra=n_extra, noise_px=1.0)
    obs = np.random.default_rng(seed).integers(0, 9, len(pts["valid"])).astype(np.int32)
    wa, wn, wf = oracle.search_local_points(views(oracle, frame=frame), projection(oracle, pr), points_view(oracle, pts),
                                            obs, 0.5, th, 0.8)
    ga, gn, gf = fe.ORBmatcher(0.8, True).SearchLocalPoints(views(fe, frame=frame), projection(fe, pr),
                                                           points_view(fe, pts), obs, th)
    assert gn == wn and np.array_equal(ga, wa) and wn > 300
    for k in wf:
        assert np.array_equal(gf[k].view(np.uint32) if gf[k].dtype == np.float32 else gf[k],
                              wf[k].view(np.uint32) if wf[k].dtype == np.float32 else wf[k]), k
    assert 0.3 < wf["in_view"].mean() < 0.98


@pytest.mark.parametrize("n_points,seed", [(1, 0), (5000, 1), (300, 2)])
def test_compute_distinctive_descriptors(oracle, n_points, seed):
    """MapPoint::ComputeDistinctiveDescriptors batched over map points."""
    from my_orb_slam2_b200 import orbfe
    desc, offsets = synth.observations_case(n_points, seed, long_lists=min(3, n_points))
    wb, wd = oracle.compute_distinctive_descriptors(desc, offsets)
    gb, gd = orbfe.Matcher().compute_distinctive_descriptors(desc, offsets)
    assert np.array_equal(gb, wb) and np.array_equal(gd, wd)
    if n_points > 100:
        assert (wb == -1).sum() > 0 and (wb > 0).sum() > n_points // 4
