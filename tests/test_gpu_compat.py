"""The C++ drop-in classes (compat/: ORB_SLAM2::ORBextractor with the reference's signature, and the
Frame::ComputeStereoMatches body) linked against liborbfe.so give the oracle's results."""
import os
import subprocess

import numpy as np
import pytest

from my_orb_slam2_b200 import synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cpp_dropin_matches_oracle(oracle, tmp_path):
    exe = os.path.join(ROOT, "compat", "_build", "test_compat")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "compat")])
    s = synth.SETTINGS["euroc"]
    w, h = synth.SHAPES["euroc"]
    left, right = synth.synth_pair(w, h, 5)
    left.tofile(tmp_path / "l.raw")
    right.tofile(tmp_path / "r.raw")
    out = tmp_path / "out.bin"
    subprocess.check_call([exe, str(tmp_path / "l.raw"), str(tmp_path / "r.raw"), str(w), str(h), str(s["nfeatures"]),
                           repr(float(np.float32(s["bf"]))), repr(float(np.float32(s["fx"]))), str(out)])
    raw = out.read_bytes()
    nl, nr, nm = np.frombuffer(raw[:12], np.int32)
    off = 12
    kp = np.frombuffer(raw[off:off + nl * 28], oracle.KEYPOINT_DTYPE); off += nl * 28
    desc = np.frombuffer(raw[off:off + nl * 32], np.uint8).reshape(nl, 32); off += nl * 32
    ur = np.frombuffer(raw[off:off + nl * 4], np.float32); off += nl * 4
    dp = np.frombuffer(raw[off:off + nl * 4], np.float32)
    params = (s["nfeatures"], s["scale_factor"], s["nlevels"], s["ini_th"], s["min_th"])
    oL, oR = oracle.Extractor(*params), oracle.Extractor(*params)
    okl, odl = oL.extract(left)
    okr, odr = oR.extract(right)
    mbf = np.float32(s["bf"])
    our, odp, on = oracle.stereo_matches(oL, oR, okl, odl, okr, odr, mbf, mbf / np.float32(s["fx"]))
    assert nl == len(okl) and nr == len(okr) and nm == on
    assert np.array_equal(kp, okl) and np.array_equal(desc, odl)
    assert np.array_equal(ur, our) and np.array_equal(dp, odp)


def test_cpp_dropin_mono_rgbd_tail_matches_oracle(oracle, tmp_path):
    """compat/FrameMono.h: the bodies of Frame::UndistortKeyPoints / ComputeImageBounds / ComputeStereoFromRGBD driven as
    the RGB-D Frame constructor does (src/Frame.cc:118-173), TUM1 camera, against the oracle."""
    exe = os.path.join(ROOT, "compat", "_build", "test_compat_mono")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "compat")])
    s = synth.SETTINGS["tum"]
    w, h = synth.SHAPES["tum"]
    img = synth.synth_image(w, h, 31)
    rng = np.random.default_rng(31)
    depth = rng.uniform(0.4, 9.0, (h, w)).astype(np.float32)
    depth[rng.random((h, w)) < 0.2] = 0.0
    img.tofile(tmp_path / "i.raw")
    depth.tofile(tmp_path / "d.f32")
    mbf = np.float32(40.0)
    out = tmp_path / "out.bin"
    subprocess.check_call([exe, str(tmp_path / "i.raw"), str(tmp_path / "d.f32"), str(w), str(h), str(s["nfeatures"]),
                           repr(float(mbf)), str(out)])
    raw = out.read_bytes()
    n = int(np.frombuffer(raw[:4], np.int32)[0])
    bounds = np.frombuffer(raw[4:20], np.float32)
    off = 20
    kp = np.frombuffer(raw[off:off + n * 28], oracle.KEYPOINT_DTYPE); off += n * 28
    kpu = np.frombuffer(raw[off:off + n * 28], oracle.KEYPOINT_DTYPE); off += n * 28
    ur = np.frombuffer(raw[off:off + n * 4], np.float32); off += n * 4
    dz = np.frombuffer(raw[off:off + n * 4], np.float32)
    cam = (517.306408, 516.469215, 318.643040, 255.313989, 0.262383, -0.953104, -0.005358, 0.002628, 1.163314)
    ok, _ = oracle.Extractor(s["nfeatures"], s["scale_factor"], s["nlevels"], s["ini_th"], s["min_th"]).extract(img)
    assert n == len(ok) and np.array_equal(kp, ok)
    oku = oracle.undistort_keypoints(ok, cam)
    assert np.array_equal(kpu, oku)
    our, odz = oracle.stereo_from_rgbd(ok, oku, depth, mbf)
    assert np.array_equal(ur, our) and np.array_equal(dz, odz)
    assert np.array_equal(bounds, oracle.image_bounds(cam, w, h))


# ---- drop-in ORBmatcher (compat/ORBmatcher.cc): the reference's C++ signatures on mock Frame / KeyFrame / MapPoint
# objects (compat/slam_min.h), results as the reference's callers observe them, against the oracle ----
def _exe():
    exe = os.path.join(ROOT, "compat", "_build", "test_compat_matcher")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "compat")])
    return exe


def _dump(d, **arrays):
    os.makedirs(d, exist_ok=True)
    for k, v in arrays.items():
        np.ascontiguousarray(v).tofile(os.path.join(d, k + ".bin"))


def _dump_frame(d, frame):
    _dump(d, keys_un=frame["keys_un"], u_right=frame["u_right"].astype(np.float32), desc=frame["desc"],
          mp_obs=frame["mp_obs"].astype(np.int32), scale_factors=frame["scale_factors"].astype(np.float32),
          bounds=np.array(frame["bounds"], np.float32))


@pytest.fixture(scope="module")
def kitti_frame(oracle):
    from cases import extract_with_oracle
    _, _, _, (kp, desc) = extract_with_oracle(oracle, "kitti", 0)
    w, h = synth.SHAPES["kitti"]
    return synth.frame_arrays(kp, desc, (0.0, 0.0, float(w), float(h)), seed=0)


@pytest.mark.parametrize("mode", ["map", "map_threads"])
def test_cpp_matcher_search_by_projection_map(oracle, kitti_frame, tmp_path, mode):
    """mode map_threads: three std::threads repeat the call with their own stack-local ORBmatcher at the same time (the
    executable exits 3 if any concurrent result differs from the one computed alone)."""
    from cases import views
    frame = kitti_frame
    mps = synth.map_points_case(frame, 20000, 1500, seed=0)
    d = str(tmp_path)
    _dump_frame(d, frame)
    _dump(d, proj_x=mps["proj_x"], proj_y=mps["proj_y"], proj_xr=mps["proj_xr"], view_cos=mps["view_cos"],
          level=mps["level"].astype(np.int32), obs=mps["obs"].astype(np.int32), in_view=mps["in_view"],
          mp_desc=mps["desc"], params=np.array([1.0, 0.8], np.float32))
    subprocess.check_call([_exe(), mode, d])
    res = np.fromfile(os.path.join(d, "result.bin"), np.int32)
    fv, mv = views(oracle, frame=frame, mps=mps)
    want, n = oracle.search_by_projection_map(fv, mv, 1.0, 0.8)
    assert res[-1] == n and n > 500
    assert np.array_equal(res[:-1], want)


def test_cpp_matcher_search_by_projection_last(oracle, kitti_frame, tmp_path):
    from cases import views
    frame = kitti_frame
    last, cam = synth.last_frame_case(frame, seed=7)
    d = str(tmp_path)
    _dump_frame(d, frame)
    camv = np.concatenate([np.asarray(cam["tcw"], np.float32).reshape(12),
                           np.array([cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["mbf"], cam["mb"]], np.float32)])
    _dump(d, cam=camv, valid=last["valid"], last_desc=last["desc"], world_pos=last["world_pos"].astype(np.float32),
          angle=last["angle"].astype(np.float32), last_tcw=np.asarray(last["tcw"], np.float32).reshape(12),
          obs=last["obs"].astype(np.int32), octave=last["octave"].astype(np.int32),
          params=np.array([7.0, 0.0, 1.0], np.float32))
    subprocess.check_call([_exe(), "last", d])
    res = np.fromfile(os.path.join(d, "result.bin"), np.int32)
    fv, lv, cv = views(oracle, frame=frame, last=last, cam=cam)
    want, n = oracle.search_by_projection_last(fv, cv, lv, 7.0, False, True)
    assert res[-1] == n and n > 300
    # -2 (assigned, then reset to NULL by the rotation histogram) is only observable where a map point existed before
    want_obs = np.where((want == -2) & (frame["mp_obs"] < 0), -1, want)
    assert np.array_equal(res[:-1], want_obs)


def test_cpp_matcher_triangulation_and_distance(oracle, kitti_frame, tmp_path):
    from cases import views
    k1, k2, F12, ep = synth.keyframe_pair_case(kitti_frame["keys_un"], kitti_frame["desc"], seed=3)
    d = str(tmp_path)
    for sub, k in (("kf1", k1), ("kf2", k2)):
        _dump(os.path.join(d, sub), keys_un=k["keys_un"], u_right=k["u_right"], desc=k["desc"], has_mp=k["has_mp"],
              node_ids=k["node_ids"].astype(np.uint32), node_offsets=k["node_offsets"].astype(np.int32),
              node_features=k["node_features"].astype(np.int32), scale_factors=k["scale_factors"].astype(np.float32),
              level_sigma2=k["level_sigma2"].astype(np.float32))
    _dump(d, geom=np.concatenate([F12.reshape(9), np.array([ep[0], ep[1], 0.0, 0.0], np.float32)]).astype(np.float32))
    subprocess.check_call([_exe(), "tri", d])
    res = np.fromfile(os.path.join(d, "result.bin"), np.int32)
    want = oracle.search_for_triangulation(views(oracle, kf=k1), views(oracle, kf=k2), F12, float(ep[0]), float(ep[1]),
                                           False, False)
    assert res[0] == len(want) and len(want) > 50
    assert np.array_equal(res[1:].reshape(-1, 2), want)
    # DescriptorDistance
    rng = np.random.default_rng(5)
    a = rng.integers(0, 256, (64, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (64, 32), dtype=np.uint8)
    _dump(d, a=a, b=b)
    subprocess.check_call([_exe(), "dist", d])
    got = np.fromfile(os.path.join(d, "result.bin"), np.int32)
    assert np.array_equal(got, np.unpackbits(a ^ b, axis=1).sum(1))


# ---- the seven members of SURVEY.md 8(f)-1 through the C++ drop-in ----
def _dump_kf(d, k):
    _dump(d, keys_un=k["keys_un"], u_right=k["u_right"], desc=k["desc"], has_mp=k["has_mp"],
          node_ids=k["node_ids"].astype(np.uint32), node_offsets=k["node_offsets"].astype(np.int32),
          node_features=k["node_features"].astype(np.int32), scale_factors=k["scale_factors"].astype(np.float32),
          level_sigma2=k["level_sigma2"].astype(np.float32))


def _dump_points(d, p):
    _dump(d, valid=p["valid"], world_pos=p["world_pos"], max_distance=p["max_distance"],
          min_distance=p["min_distance"], pt_desc=p["desc"])
    if p["normal"] is not None:
        _dump(d, normal=p["normal"])
    else:
        _dump(d, normal=np.zeros(0, np.float32))
    if p["angle"] is not None:
        _dump(d, pt_angle=p["angle"])


def _dump_proj(d, pr):
    _dump(d, proj=np.concatenate([np.asarray(pr["tcw"], np.float32).reshape(12),
                                  np.array([pr["fx"], pr["fy"], pr["cx"], pr["cy"], pr["bf"], pr["log_scale_factor"]],
                                           np.float32)]))


def _run(mode, d):
    subprocess.check_call([_exe(), mode, d])
    res = np.fromfile(os.path.join(d, "result.bin"), np.int32)
    return res[:-1], int(res[-1])


@pytest.mark.parametrize("kf_kf", [0, 1])
def test_cpp_matcher_search_by_bow(oracle, kitti_frame, tmp_path, kf_kf):
    from cases import views
    k1, k2 = synth.bow_case(kitti_frame["keys_un"], kitti_frame["desc"], seed=5)
    d = str(tmp_path)
    _dump_kf(os.path.join(d, "kf1"), k1)
    _dump_kf(os.path.join(d, "kf2"), k2)
    _dump(d, params=np.array([0.7, 1.0], np.float32))
    got, n = _run("bow%d" % kf_kf, d)
    want, wn = oracle.search_by_bow(views(oracle, kf=k1), views(oracle, kf=k2), kf_kf, 0.7, True)
    assert n == wn and wn > 300 and np.array_equal(got, want)


def test_cpp_matcher_search_for_initialization(oracle, kitti_frame, tmp_path):
    from cases import views
    f1, f2, prev = synth.initialization_case(kitti_frame["keys_un"], kitti_frame["desc"], kitti_frame["bounds"], seed=5)
    d = str(tmp_path)
    _dump_frame(os.path.join(d, "f1"), f1)
    _dump_frame(os.path.join(d, "f2"), f2)
    _dump(d, prev=prev, params=np.array([100, 0.9, 1.0], np.float32))
    got, n = _run("init", d)
    wm, wp, wn = oracle.search_for_initialization(views(oracle, frame=f1), views(oracle, frame=f2), prev, 100, 0.9, True)
    assert n == wn and wn > 100 and np.array_equal(got, wm)
    assert np.array_equal(np.fromfile(os.path.join(d, "prev_out.bin"), np.float32).reshape(-1, 2), wp)


def test_cpp_matcher_projection_overloads(oracle, kitti_frame, tmp_path):
    from cases import points_view, projection, views
    frame = dict(kitti_frame, mp_obs=synth.frame_arrays(kitti_frame["keys_un"], kitti_frame["desc"],
                                                        kitti_frame["bounds"], seed=3, taken_frac=0.15)["mp_obs"])
    # relocalisation overload
    pts, pr = synth.points3d_case(frame, seed=5)
    d = str(tmp_path / "reloc")
    _dump_frame(d, frame)
    _dump_points(d, pts)
    _dump_proj(d, pr)
    _dump(d, params=np.array([10.0, 100.0, 1.0], np.float32))
    got, n = _run("reloc", d)
    want, wn = oracle.search_by_projection_reloc(views(oracle, frame=frame), projection(oracle, pr),
                                                 points_view(oracle, pts), 10.0, 100, True)
    assert n == wn and wn > 500 and np.array_equal(got, np.where(want == -2, -1, want))
    # Sim3 overload
    pts, pr = synth.points3d_case(frame, seed=6, sim3_scale=1.05)
    d = str(tmp_path / "sim3")
    _dump_frame(d, frame)
    _dump_points(d, pts)
    _dump_proj(d, pr)
    _dump(d, params=np.array([10.0], np.float32))
    got, n = _run("sim3proj", d)
    want, wn = oracle.search_by_projection_sim3(views(oracle, frame=frame), projection(oracle, pr),
                                                points_view(oracle, pts), 10)
    assert n == wn and wn > 500 and np.array_equal(got, want)


def _fuse_observable(got, want):
    """What the keyframe / map points show after Fuse equals the oracle's bestIdx per row, except for a row that meets
    a slot whose point an earlier row already replaced (counted in nFused, no visible effect: src/ORBmatcher.cc:1011)."""
    seen = set()
    for g, w in zip(got, want):
        if g != w and not (g == -1 and w in seen):
            return False
        if w >= 0:
            seen.add(int(w))
    return True


def test_cpp_matcher_fuse(oracle, kitti_frame, tmp_path):
    from cases import points_view, projection, views
    frame = dict(kitti_frame, mp_obs=synth.frame_arrays(kitti_frame["keys_un"], kitti_frame["desc"],
                                                        kitti_frame["bounds"], seed=4, taken_frac=0.5)["mp_obs"])
    pts, pr = synth.points3d_case(frame, seed=7)
    ow = (-pr["tcw"][:, :3].T @ pr["tcw"][:, 3]).astype(np.float32)
    d = str(tmp_path / "fuse")
    _dump_frame(d, frame)
    _dump_points(d, pts)
    _dump_proj(d, pr)
    _dump(d, params=np.array([3.0], np.float32), ow=ow)
    got, n = _run("fuse", d)
    sig = (np.float32(1.0) / (frame["scale_factors"] * frame["scale_factors"])).astype(np.float32)
    want, wn = oracle.fuse(views(oracle, frame=frame), projection(oracle, pr, ow=ow), points_view(oracle, pts), 3.0, sig)
    assert n == wn and wn > 300 and _fuse_observable(got, want)
    assert (got >= 0).sum() > 0.9 * wn
    pts, pr = synth.points3d_case(frame, seed=8, sim3_scale=0.97)
    d = str(tmp_path / "fusesim3")
    _dump_frame(d, frame)
    _dump_points(d, pts)
    _dump_proj(d, pr)
    _dump(d, params=np.array([4.0], np.float32))
    got, n = _run("fusesim3", d)
    want, wn = oracle.fuse(views(oracle, frame=frame), projection(oracle, pr), points_view(oracle, pts), 4.0, None)
    assert n == wn and wn > 300 and np.array_equal(got, want)


def test_cpp_matcher_search_by_sim3(oracle, kitti_frame, tmp_path):
    from cases import points_view, sim3_pair, views
    f1, f2, p1, p2, pair = synth.sim3_pair_case(kitti_frame, seed=5)
    d = str(tmp_path)
    _dump_frame(os.path.join(d, "f1"), f1)
    _dump_frame(os.path.join(d, "f2"), f2)
    _dump_points(os.path.join(d, "p1"), p1)
    _dump_points(os.path.join(d, "p2"), p2)
    g = np.concatenate([pair["t1w"].reshape(12), pair["t2w"].reshape(12), [pair["s12"]], pair["r12"].reshape(9),
                        pair["t12"].reshape(3), [pair["fx"], pair["fy"], pair["cx"], pair["cy"], pair["log_sf1"],
                                                 pair["log_sf2"]]]).astype(np.float32)
    _dump(d, pair=g, params=np.array([7.5], np.float32))
    got, n = _run("bysim3", d)
    want, wn = oracle.search_by_sim3(views(oracle, frame=f1), views(oracle, frame=f2), points_view(oracle, p1),
                                     points_view(oracle, p2), sim3_pair(oracle, pair), 7.5)
    assert n == wn and wn > 300 and np.array_equal(got, want)


def test_cpp_vocabulary(oracle, tmp_path):
    """compat/ORBVocabulary.h: loadFromTextFile + transform into DBoW2's std::map containers."""
    voc = synth.vocabulary_case(10, 4, seed=2)
    d = str(tmp_path)
    os.makedirs(d, exist_ok=True)
    with open(os.path.join(d, "voc.txt"), "w") as f:
        f.write(synth.vocabulary_text(voc, 0, 0))
    desc = synth.descriptors_near_words(voc, 1800, seed=4)
    _dump(d, desc=desc, params=np.array([4.0], np.float32))
    subprocess.check_call([_exe(), "voc", d])
    res = np.fromfile(os.path.join(d, "result.bin"), np.int32)
    vals = np.fromfile(os.path.join(d, "bow_vals.bin"), np.float64)
    ov = oracle.Vocabulary.from_arrays(10, 4, 0, 0, voc["parent"], voc["is_leaf"], voc["desc"], voc["weight"])
    ids, wvals, fvn, fvo, fvf, _, _, _ = ov.transform(desc, 4)
    assert res[0] == ov.n_words and res[1] == len(ids) and len(ids) > 300
    assert np.array_equal(res[2:2 + len(ids)], ids.astype(np.int32)) and np.array_equal(vals, wvals)
    p = 2 + len(ids)
    assert res[p] == len(fvn)
    p += 1
    for j, nd in enumerate(fvn):
        cnt = fvo[j + 1] - fvo[j]
        assert res[p] == nd and res[p + 1] == cnt and np.array_equal(res[p + 2:p + 2 + cnt], fvf[fvo[j]:fvo[j + 1]])
        p += 2 + cnt
    assert p == len(res)
