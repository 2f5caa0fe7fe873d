"""ctypes binding of the CPU ORACLE (oracle/_ref/liborb_oracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  The product package
(my_orb_slam2_b200) never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "liborb_oracle.so")

KEYPOINT_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"),
     ("class_id", "<i4")]
)
assert KEYPOINT_DTYPE.itemsize == 28

FLAG_LIBM_SINCOS = 1
FLAG_TIE_BY_ADDRESS = 2


class ExtractorParams(C.Structure):
    _fields_ = [("nfeatures", C.c_int32), ("scale_factor", C.c_float), ("nlevels", C.c_int32),
                ("ini_th_fast", C.c_int32), ("min_th_fast", C.c_int32)]


class ExtractorInfo(C.Structure):
    _fields_ = [("nlevels", C.c_int32), ("scale_factor", C.c_float), ("scale_factors", C.c_float * 16),
                ("inv_scale_factors", C.c_float * 16), ("level_sigma2", C.c_float * 16),
                ("inv_level_sigma2", C.c_float * 16), ("features_per_level", C.c_int32 * 16),
                ("umax", C.c_int32 * 16), ("capacity", C.c_int32)]


class FrameView(C.Structure):
    _fields_ = [("n", C.c_int32), ("keys_un", C.c_void_p), ("u_right", C.c_void_p), ("desc", C.c_void_p),
                ("mp_obs", C.c_void_p), ("min_x", C.c_float), ("min_y", C.c_float), ("max_x", C.c_float),
                ("max_y", C.c_float), ("grid_inv_w", C.c_float), ("grid_inv_h", C.c_float), ("nlevels", C.c_int32),
                ("scale_factors", C.c_void_p)]


class MapPointsView(C.Structure):
    _fields_ = [("m", C.c_int32), ("proj_x", C.c_void_p), ("proj_y", C.c_void_p), ("proj_xr", C.c_void_p),
                ("level", C.c_void_p), ("view_cos", C.c_void_p), ("in_view", C.c_void_p), ("obs", C.c_void_p),
                ("desc", C.c_void_p)]


class LastFrameView(C.Structure):
    _fields_ = [("n", C.c_int32), ("valid", C.c_void_p), ("world_pos", C.c_void_p), ("obs", C.c_void_p),
                ("desc", C.c_void_p), ("octave", C.c_void_p), ("angle", C.c_void_p), ("tcw", C.c_float * 12)]


class CameraPose(C.Structure):
    _fields_ = [("tcw", C.c_float * 12), ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
                ("mbf", C.c_float), ("mb", C.c_float)]


class Points3DView(C.Structure):
    _fields_ = [("m", C.c_int32), ("valid", C.c_void_p), ("world_pos", C.c_void_p), ("normal", C.c_void_p),
                ("max_dist_inv", C.c_void_p), ("min_dist_inv", C.c_void_p), ("max_distance", C.c_void_p),
                ("desc", C.c_void_p),
                ("angle", C.c_void_p)]


class Projection(C.Structure):
    _fields_ = [("tcw", C.c_float * 12), ("is_sim3", C.c_int32), ("has_ow", C.c_int32), ("ow", C.c_float * 3),
                ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float), ("bf", C.c_float),
                ("log_scale_factor", C.c_float)]


class Sim3Pair(C.Structure):
    _fields_ = [("t1w", C.c_float * 12), ("t2w", C.c_float * 12), ("s12", C.c_float), ("r12", C.c_float * 9),
                ("t12", C.c_float * 3), ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
                ("log_scale_factor1", C.c_float), ("log_scale_factor2", C.c_float)]


class KeyFrameView(C.Structure):
    _fields_ = [("n", C.c_int32), ("keys_un", C.c_void_p), ("u_right", C.c_void_p), ("desc", C.c_void_p),
                ("has_mp", C.c_void_p), ("n_nodes", C.c_int32), ("node_ids", C.c_void_p),
                ("node_offsets", C.c_void_p), ("node_features", C.c_void_p), ("nlevels", C.c_int32),
                ("scale_factors", C.c_void_p), ("level_sigma2", C.c_void_p)]


def build(force=False):
    """Compile the oracle with its Makefile (no-op when up to date)."""
    if force or not os.path.exists(_SO):
        subprocess.check_call(["make", "-C", _HERE] + (["-B"] if force else []))
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        L.orc_fast_atan2.restype = C.c_float
        L.orc_fast_atan2.argtypes = [C.c_float, C.c_float]
        L.orc_ic_angle.restype = C.c_float
        L.orc_ic_angle.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.orc_sincos_deg.argtypes = [C.c_float, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_orb_descriptor.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_void_p]
        L.orc_undistort_keypoints.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_image_bounds.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.orc_stereo_from_rgbd.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float,
                                           C.c_void_p, C.c_void_p]
        L.orc_extractor_create.restype = C.c_void_p
        L.orc_extractor_create.argtypes = [C.POINTER(ExtractorParams), C.c_int]
        L.orc_extractor_destroy.argtypes = [C.c_void_p]
        L.orc_extractor_info.argtypes = [C.c_void_p, C.POINTER(ExtractorInfo)]
        L.orc_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_extractor_level_size.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.orc_extractor_level_ptr.restype = C.c_void_p
        L.orc_extractor_level_ptr.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int)]
        L.orc_extractor_blurred_ptr.restype = C.c_void_p
        L.orc_extractor_blurred_ptr.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int)]
        L.orc_extractor_candidates.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_extractor_level_keypoints.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.orc_distribute_octtree.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                             C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.orc_stereo_matches.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                         C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        L.orc_search_by_projection_map.argtypes = [C.POINTER(FrameView), C.POINTER(MapPointsView), C.c_float,
                                                   C.c_float, C.c_void_p]
        L.orc_search_by_projection_last.argtypes = [C.POINTER(FrameView), C.POINTER(CameraPose),
                                                    C.POINTER(LastFrameView), C.c_float, C.c_int, C.c_int, C.c_void_p]
        L.orc_search_for_triangulation.argtypes = [C.POINTER(KeyFrameView), C.POINTER(KeyFrameView), C.c_void_p,
                                                   C.c_float, C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_search_by_bow.argtypes = [C.POINTER(KeyFrameView), C.POINTER(KeyFrameView), C.c_int, C.c_float, C.c_int,
                                        C.c_void_p]
        L.orc_search_for_initialization.argtypes = [C.POINTER(FrameView), C.POINTER(FrameView), C.c_void_p, C.c_int,
                                                    C.c_float, C.c_int, C.c_void_p]
        L.orc_search_by_projection_reloc.argtypes = [C.POINTER(FrameView), C.POINTER(Projection),
                                                     C.POINTER(Points3DView), C.c_float, C.c_int, C.c_int, C.c_void_p]
        L.orc_search_by_projection_sim3.argtypes = [C.POINTER(FrameView), C.POINTER(Projection),
                                                    C.POINTER(Points3DView), C.c_int, C.c_void_p]
        L.orc_fuse.argtypes = [C.POINTER(FrameView), C.POINTER(Projection), C.POINTER(Points3DView), C.c_float,
                               C.c_void_p, C.c_void_p]
        L.orc_search_by_sim3.argtypes = [C.POINTER(FrameView), C.POINTER(FrameView), C.POINTER(Points3DView),
                                         C.POINTER(Points3DView), C.POINTER(Sim3Pair), C.c_float, C.c_void_p]
        L.orc_log_f32.restype = C.c_float
        L.orc_log_f32.argtypes = [C.c_float]
        L.orc_predict_scale.argtypes = [C.c_float, C.c_float, C.c_float, C.c_int]
        L.orc_vocabulary_create.restype = C.c_void_p
        L.orc_vocabulary_create.argtypes = [C.c_int] * 5 + [C.c_void_p] * 4
        L.orc_vocabulary_load_text.restype = C.c_void_p
        L.orc_vocabulary_load_text.argtypes = [C.c_char_p]
        L.orc_vocabulary_destroy.argtypes = [C.c_void_p]
        L.orc_vocabulary_info.argtypes = [C.c_void_p] + [C.POINTER(C.c_int)] * 6
        L.orc_vocabulary_transform.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                               C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                               C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_search_local_points.argtypes = [C.POINTER(FrameView), C.POINTER(Projection), C.POINTER(Points3DView),
                                              C.c_void_p, C.c_float, C.c_float, C.c_float] + [C.c_void_p] * 7
        L.orc_compute_distinctive_descriptors.restype = None
        L.orc_compute_distinctive_descriptors.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_cvt_gray_u8.restype = None
        L.orc_cvt_gray_u8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_remap_linear_u8.restype = None
        L.orc_remap_linear_u8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                          C.c_void_p, C.c_int]
        L.orc_hamming_bruteforce.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int64, C.c_void_p,
                                             C.c_void_p]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _u8(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    return a


# ---------------------------------------------------------------- primitives
def resize_linear(src, dw, dh):
    src = _u8(src)
    dst = np.empty((dh, dw), np.uint8)
    lib().orc_resize_linear_u8(_p(src), src.shape[1], src.shape[0], src.shape[1], _p(dst), dw, dh, dw)
    return dst


def gaussian_blur7(src):
    src = _u8(src)
    dst = np.empty_like(src)
    lib().orc_gaussian_blur7_u8(_p(src), src.shape[1], src.shape[0], src.shape[1], _p(dst), src.shape[1])
    return dst


def fast9_16(roi, threshold, nms=True):
    roi = _u8(roi)
    cap = roi.size
    xs = np.empty(cap, np.int32)
    ys = np.empty(cap, np.int32)
    sc = np.empty(cap, np.int32)
    n = lib().orc_fast9_16(_p(roi), roi.shape[1], roi.shape[0], roi.shape[1], threshold, int(nms), cap, _p(xs),
                           _p(ys), _p(sc))
    return xs[:n].copy(), ys[:n].copy(), sc[:n].copy()


def fast_atan2(y, x):
    return lib().orc_fast_atan2(float(y), float(x))


def sincos_deg(angle, use_libm=False):
    c = np.zeros(1, np.float32)
    s = np.zeros(1, np.float32)
    lib().orc_sincos_deg(float(angle), int(use_libm), _p(c), _p(s))
    return c[0], s[0]


def descriptor_distance(a, b):
    a = _u8(a)
    b = _u8(b)
    return lib().orc_descriptor_distance(_p(a), _p(b))


def ic_angle(img, x, y):
    img = _u8(img)
    return lib().orc_ic_angle(_p(img), img.shape[1], int(x), int(y))


def orb_descriptor(blurred, x, y, angle, use_libm=False):
    blurred = _u8(blurred)
    d = np.zeros(32, np.uint8)
    lib().orc_orb_descriptor(_p(blurred), blurred.shape[1], int(x), int(y), float(angle), int(use_libm), _p(d))
    return d


def distribute_octtree(xs, ys, scores, min_x, max_x, min_y, max_y, n_target, flags=0):
    xs = np.ascontiguousarray(xs, np.int32)
    ys = np.ascontiguousarray(ys, np.int32)
    sc = np.ascontiguousarray(scores, np.int32)
    kept = np.empty(max(len(xs), 1), np.int32)
    n = lib().orc_distribute_octtree(_p(xs), _p(ys), _p(sc), len(xs), min_x, max_x, min_y, max_y, n_target, flags,
                                     len(kept), _p(kept))
    if n < 0:
        raise RuntimeError("orc_distribute_octtree failed: %d" % n)
    return kept[:n].copy()


# ---------------------------------------------------------------- extractor
class Extractor:
    """CPU oracle of ORBextractor (reference: include/ORBextractor.h:45-108)."""

    def __init__(self, nfeatures, scale_factor, nlevels, ini_th, min_th, flags=0):
        self.params = ExtractorParams(nfeatures, scale_factor, nlevels, ini_th, min_th)
        self._h = lib().orc_extractor_create(C.byref(self.params), flags)
        if not self._h:
            raise ValueError("bad extractor parameters")
        self.info = ExtractorInfo()
        lib().orc_extractor_info(self._h, C.byref(self.info))
        self.nlevels = nlevels

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_extractor_destroy(self._h)
            self._h = None

    def extract(self, image):
        image = _u8(image)
        cap = self.info.capacity
        kps = np.zeros(cap, KEYPOINT_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = lib().orc_extract(self._h, _p(image), image.shape[1], image.shape[0], image.shape[1], _p(kps), _p(desc),
                              cap)
        if n < 0:
            raise RuntimeError("orc_extract failed: %d" % n)
        if n > cap:
            # the first quadtree round splits all round(w/h) root nodes before the count is looked at
            # (src/ORBextractor.cc:543-560, :574-660), so a wide image with a small per-level quota returns up to
            # 4 * round(w/h) keypoints on that level: more than info.capacity, which does not know the image size
            cap = n
            kps = np.zeros(cap, KEYPOINT_DTYPE)
            desc = np.zeros((cap, 32), np.uint8)
            n = lib().orc_extract(self._h, _p(image), image.shape[1], image.shape[0], image.shape[1], _p(kps),
                                  _p(desc), cap)
            assert n == cap
        return kps[:n].copy(), desc[:n].copy()

    def level(self, lv, blurred=False):
        w = C.c_int()
        h = C.c_int()
        if lib().orc_extractor_level_size(self._h, lv, C.byref(w), C.byref(h)) != 0:
            raise IndexError(lv)
        st = C.c_int()
        fn = lib().orc_extractor_blurred_ptr if blurred else lib().orc_extractor_level_ptr
        ptr = fn(self._h, lv, C.byref(st))
        if not ptr:
            return None
        buf = (C.c_uint8 * (w.value * h.value)).from_address(ptr)
        return np.frombuffer(buf, np.uint8).reshape(h.value, w.value).copy()

    def candidates(self, lv):
        n = lib().orc_extractor_candidates(self._h, lv, 0, None, None, None)
        xs = np.empty(n, np.int32)
        ys = np.empty(n, np.int32)
        sc = np.empty(n, np.int32)
        lib().orc_extractor_candidates(self._h, lv, n, _p(xs), _p(ys), _p(sc))
        return xs, ys, sc

    def level_keypoints(self, lv):
        n = lib().orc_extractor_level_keypoints(self._h, lv, 0, None)
        k = np.zeros(n, KEYPOINT_DTYPE)
        lib().orc_extractor_level_keypoints(self._h, lv, n, _p(k))
        return k


def stereo_matches(ext_left, ext_right, kp_l, desc_l, kp_r, desc_r, mbf, mb):
    """Frame::ComputeStereoMatches on the pyramids of the last extract() of both extractors."""
    kp_l = np.ascontiguousarray(kp_l, KEYPOINT_DTYPE)
    kp_r = np.ascontiguousarray(kp_r, KEYPOINT_DTYPE)
    desc_l = _u8(desc_l)
    desc_r = _u8(desc_r)
    ur = np.empty(len(kp_l), np.float32)
    dp = np.empty(len(kp_l), np.float32)
    n = lib().orc_stereo_matches(ext_left._h, ext_right._h, _p(kp_l), _p(desc_l), len(kp_l), _p(kp_r), _p(desc_r),
                                 len(kp_r), mbf, mb, _p(ur), _p(dp))
    return ur, dp, n


def _camera(cam):
    """(fx, fy, cx, cy, k1, k2, p1, p2[, k3]) -> the 9 floats of orbfe_camera."""
    c = np.zeros(9, np.float32)
    c[:len(cam)] = np.asarray(cam, np.float32)
    return c


def undistort_keypoints(kp, cam):
    """Frame::UndistortKeyPoints (src/Frame.cc:429-460)."""
    kp = np.ascontiguousarray(kp, KEYPOINT_DTYPE)
    out = np.empty_like(kp)
    c = _camera(cam)
    lib().orc_undistort_keypoints(_p(kp), len(kp), _p(c), _p(out))
    return out


def image_bounds(cam, cols, rows):
    """Frame::ComputeImageBounds (src/Frame.cc:462-494) -> (mnMinX, mnMaxX, mnMinY, mnMaxY)."""
    c = _camera(cam)
    b = np.zeros(4, np.float32)
    lib().orc_image_bounds(_p(c), cols, rows, _p(b))
    return b


def stereo_from_rgbd(kp, kp_un, depth, mbf):
    """Frame::ComputeStereoFromRGBD (src/Frame.cc:688-712) -> (mvuRight, mvDepth)."""
    kp = np.ascontiguousarray(kp, KEYPOINT_DTYPE)
    kp_un = np.ascontiguousarray(kp_un, KEYPOINT_DTYPE)
    depth = np.ascontiguousarray(depth, np.float32)
    ur = np.empty(len(kp), np.float32)
    dz = np.empty(len(kp), np.float32)
    lib().orc_stereo_from_rgbd(_p(kp), _p(kp_un), len(kp), _p(depth), depth.shape[1], depth.shape[0], depth.shape[1],
                               np.float32(mbf), _p(ur), _p(dz))
    return ur, dz


# ---------------------------------------------------------------- matcher views
class _Keep:
    """Holds numpy arrays alive next to the ctypes struct that points into them."""

    def __init__(self, struct, arrays):
        self.struct = struct
        self.arrays = arrays


def make_frame_view(keys_un, u_right, desc, mp_obs, bounds, scale_factors, cls=FrameView):
    keys_un = np.ascontiguousarray(keys_un, KEYPOINT_DTYPE)
    u_right = np.ascontiguousarray(u_right, np.float32)
    desc = _u8(desc)
    mp_obs = np.ascontiguousarray(mp_obs, np.int32)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    min_x, min_y, max_x, max_y = [np.float32(v) for v in bounds]
    inv_w = np.float32(64) / np.float32(max_x - min_x)  # Frame.cc:117 (static_cast<float>(COLS)/(max-min))
    inv_h = np.float32(48) / np.float32(max_y - min_y)
    v = cls(len(keys_un), _p(keys_un).value, _p(u_right).value, _p(desc).value, _p(mp_obs).value, min_x, min_y, max_x,
            max_y, inv_w, inv_h, len(sf), _p(sf).value)
    return _Keep(v, [keys_un, u_right, desc, mp_obs, sf])


def make_mappoints_view(proj_x, proj_y, proj_xr, level, view_cos, in_view, obs, desc, cls=MapPointsView):
    arrs = [np.ascontiguousarray(proj_x, np.float32), np.ascontiguousarray(proj_y, np.float32),
            np.ascontiguousarray(proj_xr, np.float32), np.ascontiguousarray(level, np.int32),
            np.ascontiguousarray(view_cos, np.float32), np.ascontiguousarray(in_view, np.uint8),
            np.ascontiguousarray(obs, np.int32), _u8(desc)]
    v = cls(len(arrs[0]), *[_p(a).value for a in arrs])
    return _Keep(v, arrs)


def make_lastframe_view(valid, world_pos, obs, desc, octave, angle, tcw, cls=LastFrameView):
    arrs = [np.ascontiguousarray(valid, np.uint8), np.ascontiguousarray(world_pos, np.float32),
            np.ascontiguousarray(obs, np.int32), _u8(desc), np.ascontiguousarray(octave, np.int32),
            np.ascontiguousarray(angle, np.float32)]
    t = (C.c_float * 12)(*[float(x) for x in np.asarray(tcw, np.float32).reshape(-1)[:12]])
    v = cls(len(arrs[0]), *[_p(a).value for a in arrs], t)
    return _Keep(v, arrs)


def make_camera_pose(tcw, fx, fy, cx, cy, mbf, mb, cls=CameraPose):
    t = (C.c_float * 12)(*[float(x) for x in np.asarray(tcw, np.float32).reshape(-1)[:12]])
    return cls(t, fx, fy, cx, cy, mbf, mb)


def make_keyframe_view(keys_un, u_right, desc, has_mp, node_ids, node_offsets, node_features, scale_factors,
                       level_sigma2, cls=KeyFrameView):
    arrs = [np.ascontiguousarray(keys_un, KEYPOINT_DTYPE), np.ascontiguousarray(u_right, np.float32), _u8(desc),
            np.ascontiguousarray(has_mp, np.uint8), np.ascontiguousarray(node_ids, np.uint32),
            np.ascontiguousarray(node_offsets, np.int32), np.ascontiguousarray(node_features, np.int32),
            np.ascontiguousarray(scale_factors, np.float32), np.ascontiguousarray(level_sigma2, np.float32)]
    v = cls(len(arrs[0]), _p(arrs[0]).value, _p(arrs[1]).value, _p(arrs[2]).value, _p(arrs[3]).value, len(arrs[4]),
            _p(arrs[4]).value, _p(arrs[5]).value, _p(arrs[6]).value, len(arrs[7]), _p(arrs[7]).value,
            _p(arrs[8]).value)
    return _Keep(v, arrs)


def search_by_projection_map(frame, mps, th, nnratio):
    assigned = np.empty(frame.struct.n, np.int32)
    n = lib().orc_search_by_projection_map(C.byref(frame.struct), C.byref(mps.struct), th, nnratio, _p(assigned))
    return assigned, n


def search_by_projection_last(cur, cam, last, th, mono, check_ori):
    assigned = np.empty(cur.struct.n, np.int32)
    n = lib().orc_search_by_projection_last(C.byref(cur.struct), C.byref(cam), C.byref(last.struct), th, int(mono),
                                            int(check_ori), _p(assigned))
    return assigned, n


def search_for_triangulation(kf1, kf2, F12, ex, ey, only_stereo, check_ori):
    F12 = np.ascontiguousarray(F12, np.float32).reshape(9)
    pairs = np.empty((max(kf1.struct.n, 1), 2), np.int32)
    n = lib().orc_search_for_triangulation(C.byref(kf1.struct), C.byref(kf2.struct), _p(F12), ex, ey,
                                           int(only_stereo), int(check_ori), _p(pairs), len(pairs))
    return pairs[:n].copy()


def hamming_bruteforce(query, mapd, index_base=0):
    query = _u8(query)
    mapd = _u8(mapd)
    best = np.empty(len(query), np.uint64)
    second = np.empty(len(query), np.uint64)
    lib().orc_hamming_bruteforce(_p(query), len(query), _p(mapd), len(mapd), index_base, _p(best), _p(second))
    return best, second


# ---------------------------------------------------------------- the seven other matcher members
def make_points3d_view(valid, world_pos, normal, max_distance, min_distance, desc, angle=None, cls=Points3DView):
    max_distance = np.ascontiguousarray(max_distance, np.float32)
    arrs = [np.ascontiguousarray(valid, np.uint8), np.ascontiguousarray(world_pos, np.float32),
            None if normal is None else np.ascontiguousarray(normal, np.float32),
            np.float32(1.2) * max_distance, np.float32(0.8) * np.ascontiguousarray(min_distance, np.float32),
            max_distance, _u8(desc),
            None if angle is None else np.ascontiguousarray(angle, np.float32)]
    v = cls(len(arrs[0]), *[None if a is None else _p(a).value for a in arrs])
    return _Keep(v, arrs)


def _fl(vals, n):
    return (C.c_float * n)(*[float(x) for x in np.asarray(vals, np.float32).reshape(-1)[:n]])


def make_projection(tcw, fx, fy, cx, cy, bf, log_scale_factor, is_sim3=False, ow=None):
    return Projection(_fl(tcw, 12), int(is_sim3), int(ow is not None), _fl(ow if ow is not None else [0, 0, 0], 3), fx,
                      fy, cx, cy, bf, log_scale_factor)


def make_sim3_pair(t1w, t2w, s12, r12, t12, fx, fy, cx, cy, log_sf1, log_sf2):
    return Sim3Pair(_fl(t1w, 12), _fl(t2w, 12), s12, _fl(r12, 9), _fl(t12, 3), fx, fy, cx, cy, log_sf1, log_sf2)


def search_by_bow(kf1, kf2, kf_kf, nnratio, check_ori):
    n_out = kf1.struct.n if kf_kf else kf2.struct.n
    matches = np.empty(max(n_out, 1), np.int32)
    n = lib().orc_search_by_bow(C.byref(kf1.struct), C.byref(kf2.struct), int(kf_kf), C.c_float(nnratio),
                                int(check_ori), _p(matches))
    return matches[:n_out], n


def search_for_initialization(f1, f2, prev_matched, window, nnratio, check_ori):
    prev = np.array(prev_matched, np.float32).reshape(-1, 2).copy()
    m12 = np.empty(max(f1.struct.n, 1), np.int32)
    n = lib().orc_search_for_initialization(C.byref(f1.struct), C.byref(f2.struct), _p(prev), int(window),
                                            C.c_float(nnratio), int(check_ori), _p(m12))
    return m12[:f1.struct.n], prev, n


def search_by_projection_reloc(cur, proj, pts, th, orb_dist, check_ori):
    assigned = np.empty(max(cur.struct.n, 1), np.int32)
    n = lib().orc_search_by_projection_reloc(C.byref(cur.struct), C.byref(proj), C.byref(pts.struct), C.c_float(th),
                                             int(orb_dist), int(check_ori), _p(assigned))
    return assigned[:cur.struct.n], n


def search_by_projection_sim3(kf, proj, pts, th):
    assigned = np.empty(max(kf.struct.n, 1), np.int32)
    n = lib().orc_search_by_projection_sim3(C.byref(kf.struct), C.byref(proj), C.byref(pts.struct), int(th),
                                            _p(assigned))
    return assigned[:kf.struct.n], n


def fuse(kf, proj, pts, th, inv_level_sigma2=None):
    best = np.empty(max(pts.struct.m, 1), np.int32)
    sig = np.zeros(16, np.float32) if inv_level_sigma2 is None else np.ascontiguousarray(inv_level_sigma2, np.float32)
    n = lib().orc_fuse(C.byref(kf.struct), C.byref(proj), C.byref(pts.struct), C.c_float(th), _p(sig), _p(best))
    return best[:pts.struct.m], n


def search_by_sim3(kf1, kf2, p1, p2, pair, th):
    m12 = np.empty(max(p1.struct.m, 1), np.int32)
    n = lib().orc_search_by_sim3(C.byref(kf1.struct), C.byref(kf2.struct), C.byref(p1.struct), C.byref(p2.struct),
                                 C.byref(pair), C.c_float(th), _p(m12))
    return m12[:p1.struct.m], n


def search_local_points(frame, proj, pts, obs, cos_limit, th, nnratio):
    m = pts.struct.m
    obs = np.ascontiguousarray(obs, np.int32)
    assigned = np.empty(max(frame.struct.n, 1), np.int32)
    out = dict(in_view=np.zeros(max(m, 1), np.uint8), proj_x=np.zeros(max(m, 1), np.float32),
               proj_y=np.zeros(max(m, 1), np.float32), proj_xr=np.zeros(max(m, 1), np.float32),
               level=np.zeros(max(m, 1), np.int32), view_cos=np.zeros(max(m, 1), np.float32))
    n = lib().orc_search_local_points(C.byref(frame.struct), C.byref(proj), C.byref(pts.struct), _p(obs), cos_limit, th,
                                      nnratio, _p(assigned), _p(out["in_view"]), _p(out["proj_x"]), _p(out["proj_y"]),
                                      _p(out["proj_xr"]), _p(out["level"]), _p(out["view_cos"]))
    return assigned[:frame.struct.n], n, {k: v[:m] for k, v in out.items()}


def compute_distinctive_descriptors(desc, offsets):
    desc = _u8(desc).reshape(-1, 32)
    offsets = np.ascontiguousarray(offsets, np.int32)
    n = len(offsets) - 1
    best = np.empty(max(n, 1), np.int32)
    bd = np.zeros((max(n, 1), 32), np.uint8)
    lib().orc_compute_distinctive_descriptors(_p(desc), _p(offsets), n, _p(best), _p(bd))
    return best[:n], bd[:n]


def cvt_gray(img, rgb=True):
    """cv::cvtColor(img, CV_{RGB,BGR}[A]2GRAY) on an H x W x {3,4} uint8 image."""
    img = _u8(img)
    h, w, ch = img.shape
    out = np.empty((h, w), np.uint8)
    lib().orc_cvt_gray_u8(_p(img), w, h, w * ch, ch, int(rgb), _p(out), w)
    return out


def remap_linear(src, mapx, mapy):
    """cv::remap(src, mapx, mapy, INTER_LINEAR) with CV_32F maps, constant border 0."""
    src = _u8(src)
    mapx = np.ascontiguousarray(mapx, np.float32)
    mapy = np.ascontiguousarray(mapy, np.float32)
    out = np.empty(mapx.shape, np.uint8)
    lib().orc_remap_linear_u8(_p(src), src.shape[1], src.shape[0], src.shape[1], _p(mapx), _p(mapy), mapx.shape[1],
                              mapx.shape[0], _p(out), mapx.shape[1])
    return out


def log_f32(x):
    return float(lib().orc_log_f32(C.c_float(x)))


def predict_scale(max_distance, dist, log_sf, nlevels):
    return int(lib().orc_predict_scale(C.c_float(max_distance), C.c_float(dist), C.c_float(log_sf), int(nlevels)))


# ---------------------------------------------------------------- DBoW2 vocabulary transform
class Vocabulary:
    """CPU oracle of the DBoW2 vocabulary tree (TemplatedVocabulary.h:1127-1262, :1338-1424)."""

    def __init__(self, handle):
        if not handle:
            raise ValueError("bad vocabulary")
        self._h = handle
        v = [C.c_int() for _ in range(6)]
        lib().orc_vocabulary_info(self._h, *[C.byref(x) for x in v])
        self.k, self.L, self.scoring, self.weighting, self.n_nodes, self.n_words = [x.value for x in v]

    @classmethod
    def from_arrays(cls, k, L, scoring, weighting, parent, is_leaf, desc, weight):
        parent = np.ascontiguousarray(parent, np.int32)
        is_leaf = np.ascontiguousarray(is_leaf, np.uint8)
        desc = _u8(desc)
        weight = np.ascontiguousarray(weight, np.float64)
        return cls(lib().orc_vocabulary_create(k, L, scoring, weighting, len(parent), _p(parent), _p(is_leaf),
                                               _p(desc), _p(weight)))

    @classmethod
    def load_text(cls, path):
        return cls(lib().orc_vocabulary_load_text(os.fsencode(path)))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_vocabulary_destroy(self._h)
            self._h = None

    def transform(self, desc, levelsup):
        desc = _u8(desc).reshape(-1, 32)
        n = len(desc)
        cap = max(n, 1)
        bow_ids, bow_vals = np.empty(cap, np.uint32), np.empty(cap, np.float64)
        fv_nodes, fv_off, fv_feat = np.empty(cap, np.uint32), np.zeros(cap + 1, np.int32), np.empty(cap, np.int32)
        word, node, wt = np.empty(cap, np.uint32), np.empty(cap, np.uint32), np.empty(cap, np.float64)
        nb, nf = C.c_int32(), C.c_int32()
        rc = lib().orc_vocabulary_transform(self._h, _p(desc), n, levelsup, _p(bow_ids), _p(bow_vals), cap, C.byref(nb),
                                            _p(fv_nodes), _p(fv_off), _p(fv_feat), cap, C.byref(nf), _p(word), _p(node),
                                            _p(wt))
        assert rc == 0
        nb, nf = nb.value, nf.value
        return (bow_ids[:nb], bow_vals[:nb], fv_nodes[:nf], fv_off[:nf + 1], fv_feat[:fv_off[nf]], word[:n], node[:n],
                wt[:n])
