"""Seeded synthetic inputs of TUM / KITTI / EuRoC shape (no datasets offline).

SURVEY.md 8(d): mid-grey + low-frequency background + filled rectangles with
random intensity steps + pixel noise, which gives a realistic FAST corner
density (~10 k candidates over 8 levels at KITTI shape).  Stereo pairs render
the same rectangle list twice, each rectangle shifted left in the right image
by its own (quarter-pixel) disparity, so Frame::ComputeStereoMatches has true
matches with sub-pixel offsets.  numpy only, deterministic for a given seed.
"""
import numpy as np

SHAPES = {
    "tum": (640, 480),     # Examples/Monocular/TUM1.yaml
    "kitti": (1241, 376),  # Examples/Stereo/KITTI00-02.yaml
    "euroc": (752, 480),   # Examples/Stereo/EuRoC.yaml
}

# (nFeatures, scaleFactor, nLevels, iniThFAST, minThFAST, fx, bf) from the reference's settings files
SETTINGS = {
    "tum": dict(nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, fx=517.306408, bf=0.0),
    "kitti": dict(nfeatures=2000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, fx=718.856, bf=386.1448),
    "euroc": dict(nfeatures=1200, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, fx=435.2046959714599,
                  bf=47.90639384423901),
}

_SS = 4  # horizontal supersampling (quarter-pixel disparities)


def _upsample_bilinear(small, h, w):
    sh, sw = small.shape
    ys = np.linspace(0, sh - 1, h)
    xs = np.linspace(0, sw - 1, w)
    y0 = np.floor(ys).astype(np.int64)
    x0 = np.floor(xs).astype(np.int64)
    y1 = np.minimum(y0 + 1, sh - 1)
    x1 = np.minimum(x0 + 1, sw - 1)
    fy = (ys - y0)[:, None]
    fx = (xs - x0)[None, :]
    top = small[y0][:, x0] * (1 - fx) + small[y0][:, x1] * fx
    bot = small[y1][:, x0] * (1 - fx) + small[y1][:, x1] * fx
    return top * (1 - fy) + bot * fy


def _scene(w, h, seed, n_rect=None):
    rng = np.random.default_rng(seed)
    if n_rect is None:
        n_rect = int(round(400 * (w * h) / (1241.0 * 376.0)))
    pad = 128  # background is rendered wider so that the right view can be a shifted crop
    bg_small = rng.normal(0.0, 1.0, (h // 6 + 2, (w + pad) // 6 + 2))
    bg = 110.0 + 12.0 * _upsample_bilinear(bg_small, h, w + pad)
    rects = []
    for _ in range(n_rect):
        rw = int(rng.integers(6, 61))
        rh = int(rng.integers(6, 61))
        x = int(rng.integers(-20, w + 20))
        y = int(rng.integers(-20, h + 20))
        step = float(rng.uniform(-90, 90))
        disp4 = int(rng.integers(2 * _SS, 96 * _SS + 1))  # disparity in quarter pixels
        rects.append((disp4, x, y, rw, rh, step))
    rects.sort(key=lambda r: r[0])  # far first, near rectangles occlude
    return bg, rects, pad


def _render(w, h, bg, rects, pad, right, noise_rng):
    bgd = 2 if right else 0
    base = bg[:, bgd:bgd + w]
    canvas = np.repeat(base, _SS, axis=1)  # h x (w*SS)
    for disp4, x, y, rw, rh, step in rects:
        x0 = x * _SS - (disp4 if right else 0)
        x1 = x0 + rw * _SS
        y0, y1 = max(y, 0), min(y + rh, h)
        x0c, x1c = max(x0, 0), min(x1, w * _SS)
        if y1 <= y0 or x1c <= x0c:
            continue
        canvas[y0:y1, x0c:x1c] = 110.0 + step
    img = canvas.reshape(h, w, _SS).mean(axis=2)
    img = img + noise_rng.normal(0.0, 2.0, (h, w))
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def synth_image(w, h, seed=0, n_rect=None):
    """One grayscale frame (H x W uint8)."""
    bg, rects, pad = _scene(w, h, seed, n_rect)
    return _render(w, h, bg, rects, pad, False, np.random.default_rng([seed, 1]))


def synth_pair(w, h, seed=0, n_rect=None):
    """Rectified stereo pair (left, right), each H x W uint8."""
    bg, rects, pad = _scene(w, h, seed, n_rect)
    left = _render(w, h, bg, rects, pad, False, np.random.default_rng([seed, 1]))
    right = _render(w, h, bg, rects, pad, True, np.random.default_rng([seed, 2]))
    return left, right


def noise_image(w, h, seed=0):
    """Stress input: uniform noise (every pixel neighbourhood is a candidate region)."""
    return np.random.default_rng([seed, 3]).integers(0, 256, (h, w), dtype=np.uint8)


def flat_image(w, h, value=128):
    """Stress input: no keypoints at all (thr-7 fallback in every cell, nkeypoints == 0 branch)."""
    return np.full((h, w), value, np.uint8)


def random_descriptors(n, seed):
    return np.random.default_rng([seed, 4]).integers(0, 256, (n, 32), dtype=np.uint8)


def flip_bits(desc, p, rng):
    """Flip each of the 256 bits of every descriptor with probability p."""
    flips = rng.random((desc.shape[0], 256)) < p
    return desc ^ np.packbits(flips, axis=1, bitorder="little")


# --------------------------------------------------------------------------------------------------------
# Matcher workloads (BASELINE.json configs 4 and 5).  Plain dicts of numpy arrays so that
# both the product views (my_orb_slam2_b200.frontend) and the test-side checker views can be built
# from the same data.
# --------------------------------------------------------------------------------------------------------
def scale_factors(nlevels=8, sf=1.2):
    out = [np.float32(1.0)]
    for _ in range(1, nlevels):
        out.append(np.float32(float(out[-1]) * float(np.float32(sf))))
    return np.array(out, np.float32)


def frame_arrays(kps, desc, bounds, seed=0, stereo_frac=0.6, taken_frac=0.05, nlevels=8):
    """Frame members read by SearchByProjection: mvKeysUn, mvuRight, mDescriptors, mvpMapPoints[i]->Observations()."""
    rng = np.random.default_rng([seed, 10])
    n = len(kps)
    u_right = np.where(rng.random(n) < stereo_frac, kps["x"] - rng.uniform(2, 60, n).astype(np.float32), -1.0)
    mp_obs = np.where(rng.random(n) < taken_frac, rng.integers(0, 4, n), -1).astype(np.int32)
    return dict(keys_un=kps, u_right=u_right.astype(np.float32), desc=desc, mp_obs=mp_obs, bounds=bounds,
                scale_factors=scale_factors(nlevels))


def map_points_case(frame, n_map=20000, n_true=1500, seed=0, bf=386.1448):
    """Local map of config 4: n_true points that really project onto frame features + distractors."""
    rng = np.random.default_rng([seed, 11])
    kps, desc = frame["keys_un"], frame["desc"]
    n = len(kps)
    min_x, min_y, max_x, max_y = frame["bounds"]
    src = rng.integers(0, n, n_true)  # with replacement: several map points compete for one feature
    t_desc = flip_bits(desc[src], 0.08, rng)
    t_x = kps["x"][src] + rng.normal(0, 1.5, n_true)
    t_y = kps["y"][src] + rng.normal(0, 1.5, n_true)
    t_lvl = np.clip(kps["octave"][src] + rng.integers(0, 2, n_true), 0, 7)
    n_d = n_map - n_true
    d_desc = rng.integers(0, 256, (n_d, 32), dtype=np.uint8)
    d_x = rng.uniform(min_x, max_x, n_d)
    d_y = rng.uniform(min_y, max_y, n_d)
    d_lvl = rng.integers(0, 8, n_d)
    perm = rng.permutation(n_map)
    proj_x = np.concatenate([t_x, d_x])[perm].astype(np.float32)
    proj_y = np.concatenate([t_y, d_y])[perm].astype(np.float32)
    level = np.concatenate([t_lvl, d_lvl])[perm].astype(np.int32)
    mdesc = np.concatenate([t_desc, d_desc])[perm]
    z = rng.uniform(5, 80, n_map)
    proj_xr = (proj_x - bf / z).astype(np.float32)
    view_cos = rng.uniform(0.9, 1.0, n_map).astype(np.float32)
    view_cos[rng.random(n_map) < 0.3] = np.float32(0.9995)  # some inside the narrow-radius branch
    in_view = (rng.random(n_map) >= 0.1).astype(np.uint8)
    obs = rng.integers(1, 12, n_map).astype(np.int32)
    obs[rng.random(n_map) < 0.02] = 0  # map points without observations do not block later ones
    return dict(proj_x=proj_x, proj_y=proj_y, proj_xr=proj_xr, level=level, view_cos=view_cos, in_view=in_view,
                obs=obs, desc=mdesc)


def last_frame_case(frame, seed=0, fx=718.856, fy=718.856, cx=607.1928, cy=185.2157, bf=386.1448, tz=0.0):
    """A previous frame whose map points re-project near the current frame's features (config 4, overload ii).
    tz: translation of the last camera along the optical axis relative to the current one (>mb: forward)."""
    rng = np.random.default_rng([seed, 12])
    kps, desc = frame["keys_un"], frame["desc"]
    n = len(kps)
    ang = 0.01
    R = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]])
    t = np.array([0.05, -0.02, 0.1])
    tcw = np.hstack([R, t[:, None]]).astype(np.float32)
    z = rng.uniform(4, 60, n)
    u = kps["x"] + rng.normal(0, 2.0, n)
    v = kps["y"] + rng.normal(0, 2.0, n)
    pc = np.stack([(u - cx) / fx * z, (v - cy) / fy * z, z], 1)
    pw = (pc - t) @ R  # R^T (pc - t)
    tl = t + np.array([0.0, 0.0, tz])
    tcw_last = np.hstack([R, tl[:, None]]).astype(np.float32)
    valid = (rng.random(n) < 0.8).astype(np.uint8)
    obs = rng.integers(1, 9, n).astype(np.int32)
    obs[rng.random(n) < 0.03] = 0
    angle = (kps["angle"] + 20.0 + rng.normal(0, 4.0, n)) % 360.0
    wild = rng.random(n) < 0.1
    angle[wild] = rng.uniform(0, 360, wild.sum())
    d = flip_bits(desc, 0.06, rng)
    octave = np.clip(kps["octave"] + rng.integers(-1, 2, n), 0, 7).astype(np.int32)
    return dict(valid=valid, world_pos=pw.astype(np.float32), obs=obs, desc=d, octave=octave,
                angle=angle.astype(np.float32), tcw=tcw_last), dict(tcw=tcw, fx=fx, fy=fy, cx=cx, cy=cy, mbf=bf,
                                                                    mb=bf / fx)


def keyframe_pair_case(kps, desc, seed=0, n_nodes=100, nlevels=8):
    """Two keyframes for SearchForTriangulation: kf2 = kf1 shifted along x (epipolar lines are image rows)."""
    rng = np.random.default_rng([seed, 13])
    n = len(kps)
    sf = scale_factors(nlevels)
    sig2 = (sf * sf).astype(np.float32)
    k2 = kps.copy()
    k2["x"] = kps["x"] - rng.uniform(3, 40, n).astype(np.float32)
    k2["y"] = kps["y"] + rng.normal(0, 0.7, n).astype(np.float32)
    k2["angle"] = (kps["angle"] + rng.normal(0, 3.0, n)) % 360.0
    d2 = flip_bits(desc, 0.05, rng)
    perm = rng.permutation(n)  # kf2 stores its features in another order
    k2, d2 = k2[perm], d2[perm]
    node1 = rng.integers(0, n_nodes, n).astype(np.uint32) * 7 + 3
    node2 = node1[perm].copy()
    moved = rng.random(n) < 0.1
    node2[moved] = rng.integers(0, n_nodes + 20, moved.sum()).astype(np.uint32) * 7 + 3

    def csr(node):
        ids = np.unique(node)
        order = np.lexsort((np.arange(len(node)), node))
        counts = np.array([(node == i).sum() for i in ids])
        return ids.astype(np.uint32), np.concatenate([[0], np.cumsum(counts)]).astype(np.int32), order.astype(np.int32)

    def kf(k, d, node, s):
        r = np.random.default_rng([seed, s])
        ids, offs, feats = csr(node)
        return dict(keys_un=k, u_right=np.where(r.random(len(k)) < 0.5, k["x"] - 5.0, -1.0).astype(np.float32), desc=d,
                    has_mp=(r.random(len(k)) < 0.3).astype(np.uint8), node_ids=ids, node_offsets=offs,
                    node_features=feats, scale_factors=sf, level_sigma2=sig2)

    F12 = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32)
    return kf(kps, desc, node1, 14), kf(k2, d2, node2, 15), F12, (np.float32(2000.0), np.float32(180.0))


def bruteforce_case(nq=8192, nm=1 << 20, n_planted=4096, seed=1):
    """Config 5: uniform random queries / map with planted near-duplicates (<= 20 flipped bits)."""
    rng = np.random.default_rng([seed, 16])
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    m = rng.integers(0, 256, (nm, 32), dtype=np.uint8)
    n_planted = min(n_planted, nq, nm)
    qi = rng.permutation(nq)[:n_planted]
    mi = rng.permutation(nm)[:n_planted]
    planted = q[qi].copy()
    for r in range(n_planted):
        bits = rng.permutation(256)[:rng.integers(0, 21)]
        for bpos in bits:
            planted[r, bpos >> 3] ^= np.uint8(1 << (bpos & 7))
    m[mi] = planted
    return q, m


# --------------------------------------------------------------------------------------------------------
# Workloads of the seven other ORBmatcher members
# --------------------------------------------------------------------------------------------------------
KITTI_CAM = dict(fx=718.856, fy=718.856, cx=607.1928, cy=185.2157, bf=386.1448)


def _small_pose(rng, rot=0.01, trans=0.1):
    """A pose close to identity: rotation about y by `rot` rad, small translation."""
    R = np.array([[np.cos(rot), 0, np.sin(rot)], [0, 1, 0], [-np.sin(rot), 0, np.cos(rot)]])
    t = rng.uniform(-trans, trans, 3)
    return R, t


def points3d_case(frame, seed=0, cam=KITTI_CAM, n_extra=400, noise_px=2.0, sim3_scale=None, nlevels=8, sf=1.2,
                  dup_frac=0.15):
    """Map points given by 3-D position that re-project near the features of `frame` (relocalisation / Sim3 /
    Fuse overloads): point i < n comes from feature i (back-projected at a random depth, descriptor = the
    feature's with flipped bits, scale-invariance range centred on the feature's octave), a `dup_frac` of the points
    are second copies aimed at the same feature (so rows compete), and n_extra distractors are random.
    Returns (points dict, projection dict)."""
    rng = np.random.default_rng([seed, 21])
    kps, desc = frame["keys_un"], frame["desc"]
    n = len(kps)
    src = np.concatenate([np.arange(n), rng.integers(0, n, int(n * dup_frac))])
    m_true = len(src)
    R, t = _small_pose(rng)
    z = rng.uniform(4, 60, m_true)
    u = kps["x"][src] + rng.normal(0, noise_px, m_true)
    v = kps["y"][src] + rng.normal(0, noise_px, m_true)
    pc = np.stack([(u - cam["cx"]) / cam["fx"] * z, (v - cam["cy"]) / cam["fy"] * z, z], 1)
    pw = (pc - t) @ R
    ow = -R.T @ t
    d3 = np.linalg.norm(pw - ow, axis=1)
    lvl = np.clip(kps["octave"][src] + rng.integers(0, 2, m_true), 0, nlevels - 1)
    max_d = d3 * (sf ** lvl) * rng.uniform(0.93, 1.0, m_true)  # PredictScale lands on lvl (or lvl-1 near the edge)
    min_d = max_d / (sf ** (nlevels - 1))
    normal = (pw - ow) / d3[:, None]
    normal += rng.normal(0, 0.15, normal.shape)
    wide = rng.random(m_true) < 0.08
    normal[wide] = rng.normal(0, 1, (wide.sum(), 3))  # some fail the 60-degree gate
    normal /= np.linalg.norm(normal, axis=1, keepdims=True)
    d = flip_bits(desc[src], 0.06, rng)
    ang = (kps["angle"][src] + 15.0 + rng.normal(0, 4.0, m_true)) % 360.0
    wild = rng.random(m_true) < 0.1
    ang[wild] = rng.uniform(0, 360, wild.sum())
    # distractors: random positions in front of the camera, random descriptors
    ze = rng.uniform(2, 80, n_extra)
    min_x, min_y, max_x, max_y = frame["bounds"]
    ue, ve = rng.uniform(min_x - 50, max_x + 50, n_extra), rng.uniform(min_y - 50, max_y + 50, n_extra)
    pce = np.stack([(ue - cam["cx"]) / cam["fx"] * ze, (ve - cam["cy"]) / cam["fy"] * ze, ze], 1)
    behind = rng.random(n_extra) < 0.1
    pce[behind, 2] *= -1.0  # negative depth
    pwe = (pce - t) @ R
    d3e = np.linalg.norm(pwe - ow, axis=1)
    max_de = d3e * rng.uniform(0.5, 4.0, n_extra)
    order = rng.permutation(m_true + n_extra)
    pts = dict(
        valid=(rng.random(m_true + n_extra) < 0.9).astype(np.uint8)[order],
        world_pos=np.concatenate([pw, pwe]).astype(np.float32)[order],
        normal=np.concatenate([normal, (pwe - ow) / d3e[:, None]]).astype(np.float32)[order],
        max_distance=np.concatenate([max_d, max_de]).astype(np.float32)[order],
        min_distance=np.concatenate([min_d, max_de / (sf ** (nlevels - 1))]).astype(np.float32)[order],
        desc=np.concatenate([d, rng.integers(0, 256, (n_extra, 32), dtype=np.uint8)])[order],
        angle=np.concatenate([ang, rng.uniform(0, 360, n_extra)]).astype(np.float32)[order])
    T = np.hstack([R, t[:, None]])
    if sim3_scale is not None:
        T = T * sim3_scale  # Scw = [s R | s t]
    proj = dict(tcw=T.astype(np.float32), fx=cam["fx"], fy=cam["fy"], cx=cam["cx"], cy=cam["cy"], bf=cam["bf"],
                log_scale_factor=float(np.log(np.float32(sf)).astype(np.float32)), is_sim3=sim3_scale is not None,
                ow=None)
    return pts, proj


def bow_case(kps, desc, seed=0, n_nodes=100, mp_frac=0.75):
    """Two feature sets with DBoW2-style feature vectors for SearchByBoW: the pair of keyframe_pair_case with a
    map-point density fit for tracking (most keyframe features carry a map point)."""
    k1, k2, _, _ = keyframe_pair_case(kps, desc, seed, n_nodes)
    r = np.random.default_rng([seed, 22])
    k1["has_mp"] = (r.random(len(k1["keys_un"])) < mp_frac).astype(np.uint8)
    k2["has_mp"] = (r.random(len(k2["keys_un"])) < mp_frac).astype(np.uint8)
    keys2 = k2["keys_un"].copy()
    wild = r.random(len(keys2)) < 0.08  # some matches must fail the rotation-consistency check
    keys2["angle"][wild] = r.uniform(0, 360, wild.sum()).astype(np.float32)
    k2["keys_un"] = keys2
    return k1, k2


def initialization_case(kps, desc, bounds, seed=0, shift=12.0, nlevels=8):
    """Two monocular frames for SearchForInitialization: F2 = F1 moved by a few pixels, other feature order, a
    fraction of near-duplicate descriptors so that later rows steal earlier rows' matches."""
    rng = np.random.default_rng([seed, 23])
    n = len(kps)
    k2 = kps.copy()
    k2["x"] = kps["x"] + shift + rng.normal(0, 3.0, n).astype(np.float32)
    k2["y"] = kps["y"] + rng.normal(0, 3.0, n).astype(np.float32)
    k2["angle"] = (kps["angle"] + rng.normal(0, 3.0, n)) % 360.0
    wild = rng.random(n) < 0.08
    k2["angle"][wild] = rng.uniform(0, 360, wild.sum())
    d2 = flip_bits(desc, 0.04, rng)
    perm = rng.permutation(n)
    k2, d2 = k2[perm], d2[perm]
    d1 = desc.copy()
    twins = rng.permutation(n)[: n // 5]  # rows that look like another row's best match
    d1[twins] = flip_bits(desc[np.roll(twins, 1)], 0.03, rng)
    k1 = kps.copy()
    k1["x"][twins] = kps["x"][np.roll(twins, 1)] + rng.normal(0, 4.0, len(twins)).astype(np.float32)
    k1["y"][twins] = kps["y"][np.roll(twins, 1)] + rng.normal(0, 4.0, len(twins)).astype(np.float32)
    none = np.full(n, -1.0, np.float32)
    f1 = dict(keys_un=k1, u_right=none, desc=d1, mp_obs=np.full(n, -1, np.int32), bounds=bounds,
              scale_factors=scale_factors(nlevels))
    f2 = dict(keys_un=k2, u_right=none, desc=d2, mp_obs=np.full(n, -1, np.int32), bounds=bounds,
              scale_factors=scale_factors(nlevels))
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)  # vbPrevMatched starts as F1's keypoints
    return f1, f2, prev


def sim3_pair_case(frame, seed=0, cam=KITTI_CAM, nlevels=8, sf=1.2, s12=1.03):
    """Two keyframes seeing the same points for SearchBySim3: KF1 = `frame`; KF2 sees every point of KF1 from another
    pose (features in another order, descriptors with flipped bits); one map point per feature on both sides."""
    rng = np.random.default_rng([seed, 24])
    kps, desc = frame["keys_un"], frame["desc"]
    n = len(kps)
    min_x, min_y, max_x, max_y = frame["bounds"]
    R1, t1 = _small_pose(rng, 0.0, 0.05)
    R2, t2 = _small_pose(rng, 0.02, 0.3)
    z = rng.uniform(6, 50, n)
    pc1 = np.stack([(kps["x"] - cam["cx"]) / cam["fx"] * z, (kps["y"] - cam["cy"]) / cam["fy"] * z, z], 1)
    pw = (pc1 - t1) @ R1
    pc2 = pw @ R2.T + t2
    u2 = cam["fx"] * pc2[:, 0] / pc2[:, 2] + cam["cx"] + rng.normal(0, 1.0, n)
    v2 = cam["fy"] * pc2[:, 1] / pc2[:, 2] + cam["cy"] + rng.normal(0, 1.0, n)
    k2 = kps.copy()
    k2["x"], k2["y"] = u2.astype(np.float32), v2.astype(np.float32)
    out = (u2 < min_x) | (u2 >= max_x) | (v2 < min_y) | (v2 >= max_y)
    k2["x"][out] = rng.uniform(min_x, max_x - 1, out.sum())
    k2["y"][out] = rng.uniform(min_y, max_y - 1, out.sum())
    d2 = flip_bits(desc, 0.05, rng)
    d2[out] = rng.integers(0, 256, (out.sum(), 32), dtype=np.uint8)
    perm = rng.permutation(n)
    k2, d2 = k2[perm], d2[perm]
    pw2 = (pw + rng.normal(0, 0.02, pw.shape))[perm]  # KF2's own map points sit (almost) on KF1's
    lvl = kps["octave"]

    def pts(world, centre, levels, s):
        r = np.random.default_rng([seed, s])
        d3 = np.linalg.norm(world - centre, axis=1)
        max_d = d3 * (sf ** levels) * r.uniform(0.93, 1.0, len(world))
        return dict(valid=(r.random(len(world)) < 0.85).astype(np.uint8), world_pos=world.astype(np.float32),
                    normal=None, max_distance=max_d.astype(np.float32),
                    min_distance=(max_d / sf ** (nlevels - 1)).astype(np.float32),
                    desc=flip_bits(desc if s == 25 else d2, 0.03, r), angle=None)

    p1 = pts(pw, -R1.T @ t1, lvl, 25)
    p2 = pts(pw2, -R2.T @ t2, lvl[perm], 26)
    f2 = dict(frame, keys_un=k2, desc=d2, u_right=np.full(n, -1.0, np.float32), mp_obs=np.full(n, -1, np.int32))
    # camera 1 from camera 2 with a scale drift: Xc1 = s12 R12 Xc2 + t12
    R12 = R1 @ R2.T
    t12 = t1 - R12 @ t2
    lsf = float(np.log(np.float32(sf)).astype(np.float32))
    pair = dict(t1w=np.hstack([R1, t1[:, None]]).astype(np.float32), t2w=np.hstack([R2, t2[:, None]]).astype(np.float32),
                s12=float(s12), r12=R12.astype(np.float32), t12=t12.astype(np.float32), fx=cam["fx"], fy=cam["fy"],
                cx=cam["cx"], cy=cam["cy"], log_sf1=lsf, log_sf2=lsf)
    return dict(frame, mp_obs=np.full(n, -1, np.int32)), f2, p1, p2, pair


# --------------------------------------------------------------------------------------------------------
# DBoW2 vocabulary: Vocabulary/ORBvoc.txt is not in the reference tree, so a random tree
# --------------------------------------------------------------------------------------------------------
def vocabulary_case(k=10, L=4, seed=0, prune=0.03, stop_frac=0.05, dup_frac=0.02):
    """A random k-ary tree of depth L in the order a DBoW2 text file lists it (parents before children): child
    descriptors are the parent's with ~20 % of the bits flipped (so the descent is meaningful), a few inner nodes are
    pruned into early leaves, a few words are stopped (weight 0) and a few siblings share a descriptor (ties).
    Returns dict(parent, is_leaf, desc, weight, k, L) with node 0 = root."""
    rng = np.random.default_rng([seed, 31])
    parent, depth, desc = [0], [0], [np.zeros(32, np.uint8)]
    frontier = [0]
    for level in range(1, L + 1):
        nxt = []
        for pnode in frontier:
            base = desc[pnode] if pnode else rng.integers(0, 256, 32, dtype=np.uint8)
            kids = flip_bits(np.repeat(base[None], k, 0), 0.5 if pnode == 0 else 0.2, rng)
            if rng.random() < dup_frac * 5:
                kids[int(rng.integers(1, k))] = kids[0]  # equal distance: the first child must win
            for c in range(k):
                parent.append(pnode)
                depth.append(level)
                desc.append(kids[c])
                if level < L and rng.random() >= prune:
                    nxt.append(len(parent) - 1)
        frontier = nxt
    n = len(parent)
    parent = np.array(parent, np.int32)
    has_child = np.zeros(n, bool)
    has_child[parent[1:]] = True
    is_leaf = (~has_child).astype(np.uint8)
    is_leaf[0] = 0
    weight = np.where(is_leaf > 0, rng.uniform(0.5, 9.0, n), 0.0)
    weight[(rng.random(n) < stop_frac) & (is_leaf > 0)] = 0.0
    return dict(parent=parent, is_leaf=is_leaf, desc=np.stack(desc), weight=weight.astype(np.float64), k=k, L=L)


def vocabulary_text(voc, scoring=0, weighting=0):
    """The text form TemplatedVocabulary::saveToTextFile writes / loadFromTextFile reads."""
    lines = ["%d %d  %d %d" % (voc["k"], voc["L"], scoring, weighting)]
    for i in range(1, len(voc["parent"])):
        lines.append("%d %d %s %s" % (voc["parent"][i], voc["is_leaf"][i], " ".join(str(int(b)) for b in voc["desc"][i]),
                                      repr(float(voc["weight"][i]))))
    return "\n".join(lines) + "\n"


def descriptors_near_words(voc, n, seed=0, flip=0.08):
    """n descriptors drawn near random leaves of the vocabulary (plus 10 % uniform ones)."""
    rng = np.random.default_rng([seed, 32])
    leaves = np.flatnonzero(voc["is_leaf"] > 0)
    pick = leaves[rng.integers(0, len(leaves), n)]
    d = flip_bits(voc["desc"][pick], flip, rng)
    uni = rng.random(n) < 0.1
    d[uni] = rng.integers(0, 256, (int(uni.sum()), 32), dtype=np.uint8)
    return d


def observations_case(n_points=5000, seed=0, max_obs=40, long_lists=3):
    """Descriptor lists of map points for ComputeDistinctiveDescriptors: every point has 0..max_obs noisy copies of a
    prototype (a few outliers included), and `long_lists` points carry several hundred observations."""
    rng = np.random.default_rng([seed, 41])
    counts = rng.integers(0, max_obs + 1, n_points)
    counts[rng.random(n_points) < 0.3] = rng.integers(1, 5, int((rng.random(n_points) < 0.3).sum() or 1))[0]
    counts[:long_lists] = rng.integers(100, 400, long_lists)
    offsets = np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
    proto = rng.integers(0, 256, (n_points, 32), dtype=np.uint8)
    desc = flip_bits(np.repeat(proto, counts, axis=0), 0.07, rng)
    out = rng.random(len(desc)) < 0.05
    desc[out] = rng.integers(0, 256, (int(out.sum()), 32), dtype=np.uint8)
    return desc, offsets
