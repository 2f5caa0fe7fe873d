"""ctypes binding of liborbfe.so — the C ABI declared in include/orbfe.h.

The library is the product: hand-written sm_100a CUDA kernels behind plain-C entry points.  This module only
marshals numpy arrays; it never computes results itself and has NO fallback: if the shared library is missing
or no CUDA device works, the calls raise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liborbfe.so")

KEYPOINT_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"),
     ("class_id", "<i4")]
)
assert KEYPOINT_DTYPE.itemsize == 28  # layout of cv::KeyPoint / orbfe_keypoint


ERR_ARG, ERR_CUDA, ERR_CAPACITY, ERR_STATE = -1, -2, -3, -4  # include/orbfe.h:30-34


class OrbfeError(RuntimeError):
    """A non-zero status of the C ABI; `code` is the ORBFE_ERR_* value (None when the library itself is missing)."""

    def __init__(self, msg, code=None):
        super().__init__(msg)
        self.code = code


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


class KeyFrameView(C.Structure):
    _fields_ = [("n", C.c_int32), ("keys_un", C.c_void_p), ("u_right", C.c_void_p), ("desc", C.c_void_p),
                ("has_mp", C.c_void_p), ("n_nodes", C.c_int32), ("node_ids", C.c_void_p),
                ("node_offsets", C.c_void_p), ("node_features", C.c_void_p), ("nlevels", C.c_int32),
                ("scale_factors", C.c_void_p), ("level_sigma2", C.c_void_p)]


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


# every symbol include/orbfe.h declares: name -> (restype, argtypes)
_P = C.c_void_p
_I = C.c_int
_F = C.c_float
SYMBOLS = {
    "orbfe_last_error": (C.c_char_p, []),
    "orbfe_version": (C.c_char_p, []),
    "orbfe_device_count": (_I, []),
    "orbfe_extractor_create": (_I, [C.POINTER(ExtractorParams), _I, _I, _I, _I, C.POINTER(_P)]),
    "orbfe_extractor_destroy": (None, [_P]),
    "orbfe_extractor_get_info": (_I, [_P, C.POINTER(ExtractorInfo)]),
    "orbfe_extract": (_I, [_P, _P, _I, _P, _P, _I, _P]),
    "orbfe_extract_batch": (_I, [_P, _P, C.c_size_t, _I, _I, _P, _P, _I, _P]),
    "orbfe_extract_upload": (_I, [_P, _P, C.c_size_t, _I, _I, _I]),
    "orbfe_extract_run": (_I, [_P, _I]),
    "orbfe_extract_upload_color": (_I, [_P, _P, C.c_size_t, _I, _I, _I, _I]),
    "orbfe_extractor_set_rectify_maps": (_I, [_P, _P, _P, _I, _I]),
    "orbfe_extract_upload_distorted": (_I, [_P, _P, C.c_size_t, _I, _I]),
    "orbfe_extract_download": (_I, [_P, _I, _P, _P, _I, _P]),
    "orbfe_extractor_sync": (_I, [_P]),
    "orbfe_extract_download_async": (_I, [_P, _I, _P, _P, _I, _P]),
    "orbfe_extract_finish": (_I, [_P]),
    "orbfe_stereo_match_download_async": (_I, [_P, _I, _P, _P, _I, _P]),
    "orbfe_extractor_share_stream": (_I, [_P, _P]),
    "orbfe_extractor_stream": (_P, [_P]),
    "orbfe_extractor_launch_count": (C.c_int64, [_P]),
    "orbfe_extractor_set_profiling": (_I, [_P, _I]),
    "orbfe_extractor_stage_times": (_I, [_P, _P, _P, _P]),
    "orbfe_extractor_level": (_I, [_P, _I, _I, C.POINTER(_P), C.POINTER(_I), C.POINTER(_I), C.POINTER(_I)]),
    "orbfe_extractor_download_level": (_I, [_P, _I, _I, _P, _I]),
    "orbfe_extractor_device_results": (_I, [_P, C.POINTER(_P), C.POINTER(_P), C.POINTER(_P), C.POINTER(_I)]),
    "orbfe_extractor_debug_candidates": (_I, [_P, _I, _I, _I, _P, _P, _P, _P]),
    "orbfe_extractor_debug_level_keypoints": (_I, [_P, _I, _I, _I, _P, _P, _P, _P]),
    "orbfe_stereo_match": (_I, [_P, _P, _I, _F, _F, _P, _P, _I, _P]),
    "orbfe_stereo_frame": (_I, [_P, _P, _P, _P, C.c_size_t, _I, _I, _F, _F, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I]),
    "orbfe_stereo_match_run": (_I, [_P, _P, _I, _F, _F]),
    "orbfe_stereo_match_download": (_I, [_P, _I, _P, _P, _I, _P]),
    "orbfe_undistort_keypoints": (_I, [_P, _I, _P, _P, _I]),
    "orbfe_image_bounds": (_I, [_P, _I, _I, _P]),
    "orbfe_stereo_from_rgbd": (_I, [_P, _I, _P, C.c_size_t, _I, _F, _P, _P, _I]),
    "orbfe_matcher_create": (_I, [_I, C.POINTER(_P)]),
    "orbfe_matcher_destroy": (None, [_P]),
    "orbfe_matcher_launch_count": (C.c_int64, [_P]),
    "orbfe_matcher_stream": (_P, [_P]),
    "orbfe_descriptor_distance": (_I, [_P, _P, _P, _I, _P]),
    "orbfe_search_by_projection_map": (_I, [_P, C.POINTER(FrameView), C.POINTER(MapPointsView), _F, _F, _P, _P]),
    "orbfe_search_by_projection_last": (_I, [_P, C.POINTER(FrameView), C.POINTER(CameraPose),
                                             C.POINTER(LastFrameView), _F, _I, _I, _P, _P]),
    "orbfe_search_for_triangulation": (_I, [_P, C.POINTER(KeyFrameView), C.POINTER(KeyFrameView), _P, _F, _F, _I, _I,
                                            _P, _I, _P]),
    "orbfe_search_by_bow": (_I, [_P, C.POINTER(KeyFrameView), C.POINTER(KeyFrameView), _I, _F, _I, _P, _P]),
    "orbfe_search_for_initialization": (_I, [_P, C.POINTER(FrameView), C.POINTER(FrameView), _P, _I, _F, _I, _P, _P]),
    "orbfe_search_by_projection_reloc": (_I, [_P, C.POINTER(FrameView), C.POINTER(Projection),
                                              C.POINTER(Points3DView), _F, _I, _I, _P, _P]),
    "orbfe_search_by_projection_sim3": (_I, [_P, C.POINTER(FrameView), C.POINTER(Projection),
                                             C.POINTER(Points3DView), _I, _P, _P]),
    "orbfe_search_local_points": (_I, [_P, C.POINTER(FrameView), C.POINTER(Projection), C.POINTER(Points3DView), _P,
                                       _F, _F, _F, _P, _P, _P, _P, _P, _P, _P, _P]),
    "orbfe_compute_distinctive_descriptors": (_I, [_P, _P, _P, _I, _P, _P]),
    "orbfe_fuse": (_I, [_P, C.POINTER(FrameView), C.POINTER(Projection), C.POINTER(Points3DView), _F, _P, _P, _P]),
    "orbfe_search_by_sim3": (_I, [_P, C.POINTER(FrameView), C.POINTER(FrameView), C.POINTER(Points3DView),
                                  C.POINTER(Points3DView), C.POINTER(Sim3Pair), _F, _P, _P]),
    "orbfe_vocabulary_create": (_I, [_I, _I, _I, _I, _I, _I, _P, _P, _P, _P, C.POINTER(_P)]),
    "orbfe_vocabulary_load_text": (_I, [_I, C.c_char_p, C.POINTER(_P)]),
    "orbfe_vocabulary_destroy": (None, [_P]),
    "orbfe_vocabulary_info": (_I, [_P, _P]),
    "orbfe_vocabulary_transform": (_I, [_P, _P, _I, _I, _P, _P, _I, _P, _P, _P, _P, _I, _P, _P, _P, _P]),
    "orbfe_vocabulary_transform_device": (_I, [_P, _P, _I, _I, _P, _P, _I, _P, _P, _P, _P, _I, _P]),
    "orbfe_hamming_bruteforce": (_I, [_P, _P, _I, _P, _I, C.c_int64, _P, _P]),
    "orbfe_hamming_bruteforce_device": (_I, [_P, _P, _I, _P, _I, C.c_int64, _P, _P]),
    "orbfe_hamming_merge_device": (_I, [_P, _P, _P, _I, _I, _P, _P]),
    "orbfe_hamming_bruteforce_tc_device": (_I, [_P, _P, _I, _P, _I, C.c_int64, _P, _P]),
    "orbfe_hamming_bruteforce_device_noalloc": (_I, [_P, _P, _I, _P, _I, C.c_int64, _P, _P]),
    "orbfe_matcher_sync": (_I, [_P]),
    "orbfe_comm_unique_id": (_I, [_P]),
    "orbfe_comm_create": (_I, [_P, _I, _I, _I, _I, C.POINTER(_P)]),
    "orbfe_comm_destroy": (None, [_P]),
    "orbfe_comm_info": (_I, [_P, C.POINTER(_I), C.POINTER(_I), C.POINTER(_I), C.POINTER(_I)]),
    "orbfe_comm_status": (_I, [_P]),
    "orbfe_hamming_bruteforce_sharded": (_I, [_P, _P, _P, _I, _P, _I, C.c_int64, _P, _P, _I]),
    "orbfe_comm_set_timing": (_I, [_P, _I]),
    "orbfe_comm_last_times": (_I, [_P, C.POINTER(_F), C.POINTER(_F)]),
    "orbfe_measure_popc_peak": (_I, [_P, C.POINTER(C.c_double)]),
}

_lib = None


def lib():
    """Load liborbfe.so (fails loudly when it was not built: there is no other implementation)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise OrbfeError("%s is missing: build it with `make -C my_orb_slam2_b200/csrc` (or "
                             "__graft_entry__.build()); there is no CPU fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise OrbfeError("orbfe status %d: %s" % (rc, lib().orbfe_last_error().decode()), rc)


def device_count():
    return lib().orbfe_device_count()


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _u8(a):
    return np.ascontiguousarray(a, dtype=np.uint8)


class Extractor:
    """Thin object wrapper of the orbfe_extractor_* functions (one handle = one stream + its device arenas)."""

    def __init__(self, nfeatures, scale_factor, nlevels, ini_th, min_th, width, height, max_batch=1, device=0):
        self.params = ExtractorParams(nfeatures, scale_factor, nlevels, ini_th, min_th)
        self.width, self.height, self.max_batch, self.device = width, height, max_batch, device
        h = C.c_void_p()
        check(lib().orbfe_extractor_create(C.byref(self.params), width, height, max_batch, device, C.byref(h)))
        self._h = h
        self.info = ExtractorInfo()
        check(lib().orbfe_extractor_get_info(self._h, C.byref(self.info)))
        self.capacity = self.info.capacity
        self.nlevels = nlevels

    def close(self):
        if getattr(self, "_h", None):
            lib().orbfe_extractor_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- synchronous calls ---------------------------------------------------------------
    def extract(self, image):
        """ORBextractor::operator() on one H x W uint8 image -> (keypoints[n], descriptors[n,32])."""
        kps, desc, counts = self.extract_batch(image[None])
        return kps[0][:counts[0]].copy(), desc[0][:counts[0]].copy()

    def extract_batch(self, images):
        images = _u8(images)
        b, h, w = images.shape
        if (w, h) != (self.width, self.height):
            raise ValueError("image size %dx%d != handle size %dx%d" % (w, h, self.width, self.height))
        kps = np.zeros((b, self.capacity), KEYPOINT_DTYPE)
        desc = np.zeros((b, self.capacity, 32), np.uint8)
        counts = np.zeros(b, np.int32)
        check(lib().orbfe_extract_batch(self._h, _p(images), h * w, w, b, _p(kps), _p(desc), self.capacity,
                                        _p(counts)))
        return kps, desc, counts

    # -- asynchronous pieces ---------------------------------------------------------------
    def upload(self, images_ptr, batch, image_stride=None, row_stride=None, on_device=False):
        check(lib().orbfe_extract_upload(self._h, images_ptr, image_stride or self.width * self.height,
                                         row_stride or self.width, batch, int(on_device)))

    def upload_color(self, images, rgb=True):
        """cvtColor(*2GRAY) fused into the upload: images = [batch, H, W, 3|4] uint8 (Tracking::GrabImage*)."""
        images = _u8(images)
        b, h, w, ch = images.shape
        if (w, h) != (self.width, self.height):
            raise ValueError("image size %dx%d != handle size %dx%d" % (w, h, self.width, self.height))
        check(lib().orbfe_extract_upload_color(self._h, _p(images), h * w * ch, w * ch, ch, int(rgb), b))
        self.sync()  # the host array may go away

    def set_rectify_maps(self, map_x, map_y, src_width, src_height):
        map_x = np.ascontiguousarray(map_x, np.float32)
        map_y = np.ascontiguousarray(map_y, np.float32)
        if map_x.shape != (self.height, self.width) or map_y.shape != map_x.shape:
            raise ValueError("maps must be %d x %d" % (self.height, self.width))
        check(lib().orbfe_extractor_set_rectify_maps(self._h, _p(map_x), _p(map_y), src_width, src_height))

    def upload_distorted(self, images):
        """cv::remap(INTER_LINEAR) fused into the upload: images = [batch, srcH, srcW] uint8 raw camera frames."""
        images = _u8(images)
        b, h, w = images.shape
        check(lib().orbfe_extract_upload_distorted(self._h, _p(images), h * w, w, b))
        self.sync()

    def run(self, batch):
        check(lib().orbfe_extract_run(self._h, batch))

    def download(self, batch, kps, desc, counts):
        check(lib().orbfe_extract_download(self._h, batch, _p(kps), _p(desc), kps.shape[-1], _p(counts)))

    def sync(self):
        check(lib().orbfe_extractor_sync(self._h))

    # -- the rest of the mono / RGB-D Frame constructors (src/Frame.cc:118-228) -----------------------
    def undistort_keypoints(self, batch, cam):
        """Frame::UndistortKeyPoints on the last run's keypoints: cam = (fx, fy, cx, cy, k1, k2, p1, p2[, k3]).
        Returns mvKeysUn as [batch][capacity] keypoints (rows beyond the image's count are untouched zeros)."""
        out = np.zeros((batch, self.capacity), KEYPOINT_DTYPE)
        check(lib().orbfe_undistort_keypoints(self._h, batch, _p(camera(cam)), _p(out), self.capacity))
        return out

    def stereo_from_rgbd(self, batch, depth, mbf):
        """Frame::ComputeStereoFromRGBD after undistort_keypoints: depth [batch, H, W] float32 -> (mvuRight, mvDepth)."""
        depth = np.ascontiguousarray(depth, np.float32)
        if depth.ndim == 2:
            depth = depth[None]
        b, h, w = depth.shape
        ur = np.empty((batch, self.capacity), np.float32)
        dz = np.empty((batch, self.capacity), np.float32)
        check(lib().orbfe_stereo_from_rgbd(self._h, batch, _p(depth), h * w, w, np.float32(mbf), _p(ur), _p(dz),
                                           self.capacity))
        return ur, dz

    def share_stream(self, other):
        check(lib().orbfe_extractor_share_stream(self._h, other._h))

    @property
    def stream(self):
        return lib().orbfe_extractor_stream(self._h)

    @property
    def launch_count(self):
        return lib().orbfe_extractor_launch_count(self._h)

    def set_profiling(self, enable=True):
        check(lib().orbfe_extractor_set_profiling(self._h, int(enable)))

    def stage_times(self):
        """Summed CUDA-event milliseconds per stage since set_profiling(True): dict + number of runs covered."""
        ms = np.zeros(7, np.float32)
        runs, sruns = C.c_int32(), C.c_int32()
        check(lib().orbfe_extractor_stage_times(self._h, _p(ms), C.byref(runs), C.byref(sruns)))
        names = ["pyramid", "fast", "quadtree", "blur", "describe", "stereo_match", "stereo_median"]
        return dict(zip(names, [float(x) for x in ms])), runs.value, sruns.value

    # -- pyramid / stage dumps ------------------------------------------------------------
    def level_shape(self, lv):
        w, h, pitch, ptr = C.c_int(), C.c_int(), C.c_int(), C.c_void_p()
        check(lib().orbfe_extractor_level(self._h, 0, lv, C.byref(ptr), C.byref(pitch), C.byref(w), C.byref(h)))
        return h.value, w.value

    def level(self, lv, image_index=0):
        """mvImagePyramid[lv] of image `image_index` of the last run, copied to the host."""
        h, w = self.level_shape(lv)
        out = np.empty((h, w), np.uint8)
        check(lib().orbfe_extractor_download_level(self._h, image_index, lv, _p(out), w))
        return out

    def _debug(self, fn, lv, image_index):
        n = C.c_int32()
        check(fn(self._h, image_index, lv, 0, None, None, None, C.byref(n)))
        xs, ys, sc = (np.empty(max(n.value, 1), np.int32) for _ in range(3))
        check(fn(self._h, image_index, lv, n.value, _p(xs), _p(ys), _p(sc), C.byref(n)))
        return xs[:n.value], ys[:n.value], sc[:n.value]

    def candidates(self, lv, image_index=0):
        return self._debug(lib().orbfe_extractor_debug_candidates, lv, image_index)

    def level_keypoints(self, lv, image_index=0):
        return self._debug(lib().orbfe_extractor_debug_level_keypoints, lv, image_index)


def camera(cam):
    """(fx, fy, cx, cy, k1, k2, p1, p2[, k3]) -> the 9 floats of orbfe_camera."""
    c = np.zeros(9, np.float32)
    c[:len(cam)] = np.asarray(cam, np.float32)
    return c


def image_bounds(cam, width, height):
    """Frame::ComputeImageBounds -> (mnMinX, mnMaxX, mnMinY, mnMaxY)."""
    b = np.zeros(4, np.float32)
    check(lib().orbfe_image_bounds(_p(camera(cam)), width, height, _p(b)))
    return b


def stereo_match(left, right, batch, mbf, mb):
    """Frame::ComputeStereoMatches on the resident results of two Extractor handles -> (uRight, depth, n)."""
    ur = np.empty((batch, left.capacity), np.float32)
    dp = np.empty((batch, left.capacity), np.float32)
    n = np.zeros(batch, np.int32)
    check(lib().orbfe_stereo_match(left._h, right._h, batch, mbf, mb, _p(ur), _p(dp), left.capacity, _p(n)))
    return ur, dp, n


def stereo_frame(left, right, left_images, right_images, mbf, mb):
    """The stereo Frame constructor in one C-ABI call (orbfe_stereo_frame): (B, H, W) u8 host images of both cameras
    -> (kpL, descL, nL, kpR, descR, nR, uRight, depth, n_matched), arrays with `capacity` rows per image."""
    li = np.ascontiguousarray(left_images, np.uint8)
    ri = np.ascontiguousarray(right_images, np.uint8)
    if li.ndim == 2:
        li, ri = li[None], ri[None]
    b, h, w = li.shape
    cap = left.capacity
    kl = np.zeros((b, cap), KEYPOINT_DTYPE)
    kr = np.zeros((b, cap), KEYPOINT_DTYPE)
    dl = np.zeros((b, cap, 32), np.uint8)
    dr = np.zeros((b, cap, 32), np.uint8)
    nl, nr, nm = np.zeros(b, np.int32), np.zeros(b, np.int32), np.zeros(b, np.int32)
    ur = np.empty((b, cap), np.float32)
    dp = np.empty((b, cap), np.float32)
    check(lib().orbfe_stereo_frame(left._h, right._h, _p(li), _p(ri), h * w, w, b, mbf, mb, _p(kl), _p(dl), _p(nl),
                                   _p(kr), _p(dr), _p(nr), _p(ur), _p(dp), _p(nm), cap))
    return kl, dl, nl, kr, dr, nr, ur, dp, nm


class Matcher:
    """Thin object wrapper of the orbfe_matcher_* / search functions."""

    def __init__(self, device=0):
        h = C.c_void_p()
        check(lib().orbfe_matcher_create(device, C.byref(h)))
        self._h = h
        self.device = device

    def close(self):
        if getattr(self, "_h", None):
            lib().orbfe_matcher_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def launch_count(self):
        return lib().orbfe_matcher_launch_count(self._h)

    @property
    def stream(self):
        return lib().orbfe_matcher_stream(self._h)

    def sync(self):
        check(lib().orbfe_matcher_sync(self._h))

    def descriptor_distance(self, a, b):
        a, b = _u8(a).reshape(-1, 32), _u8(b).reshape(-1, 32)
        out = np.empty(len(a), np.int32)
        check(lib().orbfe_descriptor_distance(self._h, _p(a), _p(b), len(a), _p(out)))
        return out

    def search_by_projection_map(self, frame, mps, th, nnratio):
        assigned = np.empty(max(frame.struct.n, 1), np.int32)
        n = C.c_int32()
        check(lib().orbfe_search_by_projection_map(self._h, C.byref(frame.struct), C.byref(mps.struct), th, nnratio,
                                                   _p(assigned), C.byref(n)))
        return assigned[:frame.struct.n], n.value

    def search_by_projection_last(self, cur, cam, last, th, mono, check_ori):
        assigned = np.empty(max(cur.struct.n, 1), np.int32)
        n = C.c_int32()
        check(lib().orbfe_search_by_projection_last(self._h, C.byref(cur.struct), C.byref(cam), C.byref(last.struct),
                                                    th, int(mono), int(check_ori), _p(assigned), C.byref(n)))
        return assigned[:cur.struct.n], n.value

    def search_for_triangulation(self, kf1, kf2, F12, ex, ey, only_stereo, check_ori):
        F12 = np.ascontiguousarray(F12, np.float32).reshape(9)
        pairs = np.empty((max(kf1.struct.n, 1), 2), np.int32)
        n = C.c_int32()
        check(lib().orbfe_search_for_triangulation(self._h, C.byref(kf1.struct), C.byref(kf2.struct), _p(F12), ex, ey,
                                                   int(only_stereo), int(check_ori), _p(pairs), len(pairs),
                                                   C.byref(n)))
        return pairs[:n.value].copy()

    def search_by_bow(self, kf1, kf2, kf_kf, nnratio, check_ori):
        n_out = kf1.struct.n if kf_kf else kf2.struct.n
        matches = np.empty(max(n_out, 1), np.int32)
        n = C.c_int32()
        check(lib().orbfe_search_by_bow(self._h, C.byref(kf1.struct), C.byref(kf2.struct), int(kf_kf), nnratio,
                                        int(check_ori), _p(matches), C.byref(n)))
        return matches[:n_out], n.value

    def search_for_initialization(self, f1, f2, prev_matched, window, nnratio, check_ori):
        prev = np.array(prev_matched, np.float32).reshape(-1, 2).copy()
        assert len(prev) == f1.struct.n
        m12 = np.empty(max(f1.struct.n, 1), np.int32)
        n = C.c_int32()
        check(lib().orbfe_search_for_initialization(self._h, C.byref(f1.struct), C.byref(f2.struct), _p(prev),
                                                    int(window), nnratio, int(check_ori), _p(m12), C.byref(n)))
        return m12[:f1.struct.n], prev, n.value

    def search_by_projection_reloc(self, cur, proj, pts, th, orb_dist, check_ori):
        assigned = np.empty(max(cur.struct.n, 1), np.int32)
        n = C.c_int32()
        check(lib().orbfe_search_by_projection_reloc(self._h, C.byref(cur.struct), C.byref(proj), C.byref(pts.struct),
                                                     th, int(orb_dist), int(check_ori), _p(assigned), C.byref(n)))
        return assigned[:cur.struct.n], n.value

    def search_by_projection_sim3(self, kf, proj, pts, th):
        assigned = np.empty(max(kf.struct.n, 1), np.int32)
        n = C.c_int32()
        check(lib().orbfe_search_by_projection_sim3(self._h, C.byref(kf.struct), C.byref(proj), C.byref(pts.struct),
                                                    int(th), _p(assigned), C.byref(n)))
        return assigned[:kf.struct.n], n.value

    def search_local_points(self, frame, proj, pts, obs, cos_limit, th, nnratio):
        """Returns (assigned, nmatches, dict(in_view, proj_x, proj_y, proj_xr, level, view_cos))."""
        m = pts.struct.m
        obs = np.ascontiguousarray(obs, np.int32)
        assigned = np.empty(max(frame.struct.n, 1), np.int32)
        out = dict(in_view=np.zeros(max(m, 1), np.uint8), proj_x=np.zeros(max(m, 1), np.float32),
                   proj_y=np.zeros(max(m, 1), np.float32), proj_xr=np.zeros(max(m, 1), np.float32),
                   level=np.zeros(max(m, 1), np.int32), view_cos=np.zeros(max(m, 1), np.float32))
        n = C.c_int32()
        check(lib().orbfe_search_local_points(self._h, C.byref(frame.struct), C.byref(proj), C.byref(pts.struct),
                                              _p(obs), cos_limit, th, nnratio, _p(assigned), C.byref(n),
                                              _p(out["in_view"]), _p(out["proj_x"]), _p(out["proj_y"]),
                                              _p(out["proj_xr"]), _p(out["level"]), _p(out["view_cos"])))
        return assigned[:frame.struct.n], n.value, {k: v[:m] for k, v in out.items()}

    def fuse(self, kf, proj, pts, th, inv_level_sigma2=None):
        best = np.empty(max(pts.struct.m, 1), np.int32)
        n = C.c_int32()
        sig = None if inv_level_sigma2 is None else np.ascontiguousarray(inv_level_sigma2, np.float32)
        check(lib().orbfe_fuse(self._h, C.byref(kf.struct), C.byref(proj), C.byref(pts.struct), th,
                               None if sig is None else _p(sig), _p(best), C.byref(n)))
        return best[:pts.struct.m], n.value

    def search_by_sim3(self, kf1, kf2, p1, p2, pair, th):
        m12 = np.empty(max(p1.struct.m, 1), np.int32)
        n = C.c_int32()
        check(lib().orbfe_search_by_sim3(self._h, C.byref(kf1.struct), C.byref(kf2.struct), C.byref(p1.struct),
                                         C.byref(p2.struct), C.byref(pair), th, _p(m12), C.byref(n)))
        return m12[:p1.struct.m], n.value

    def compute_distinctive_descriptors(self, desc, offsets):
        desc = _u8(desc).reshape(-1, 32)
        offsets = np.ascontiguousarray(offsets, np.int32)
        n = len(offsets) - 1
        best = np.empty(max(n, 1), np.int32)
        bd = np.zeros((max(n, 1), 32), np.uint8)
        check(lib().orbfe_compute_distinctive_descriptors(self._h, _p(desc), _p(offsets), n, _p(best), _p(bd)))
        return best[:n], bd[:n]

    def popc_peak(self):
        v = C.c_double()
        check(lib().orbfe_measure_popc_peak(self._h, C.byref(v)))
        return v.value

    def hamming_bruteforce(self, query, mapd, index_base=0):
        query, mapd = _u8(query).reshape(-1, 32), _u8(mapd).reshape(-1, 32)
        best = np.empty(len(query), np.uint64)
        second = np.empty(len(query), np.uint64)
        check(lib().orbfe_hamming_bruteforce(self._h, _p(query), len(query), _p(mapd), len(mapd), index_base,
                                             _p(best), _p(second)))
        return best, second

    def hamming_bruteforce_device(self, d_query, nq, d_map, nm, index_base, d_best, d_second):
        """Device-pointer variant (ints from torch .data_ptr()); asynchronous on the matcher's stream."""
        check(lib().orbfe_hamming_bruteforce_device(self._h, d_query, nq, d_map, nm, index_base, d_best, d_second))

    def hamming_bruteforce_tc_device(self, d_query, nq, d_map, nm, index_base, d_best, d_second):
        """The tensor-core variant (tcgen05 kind::i8 GEMM on the expanded bits); same keys, asynchronous."""
        check(lib().orbfe_hamming_bruteforce_tc_device(self._h, d_query, nq, d_map, nm, index_base, d_best, d_second))

    def hamming_merge_device(self, d_best_all, d_second_all, shards, nq, d_best, d_second):
        check(lib().orbfe_hamming_merge_device(self._h, d_best_all, d_second_all, shards, nq, d_best, d_second))


COMM_ID_BYTES = 128
EXCHANGE_AUTO, EXCHANGE_NCCL, EXCHANGE_PEER = 0, 1, 2
KERNEL_TENSOR = 16   # OR into `exchange`: the shard's top-2 by the tensor-core kernel


def comm_unique_id():
    """Rank 0's half of the bootstrap (ncclGetUniqueId through the library): 128 bytes to hand to every rank."""
    buf = np.zeros(COMM_ID_BYTES, np.uint8)
    check(lib().orbfe_comm_unique_id(_p(buf)))
    return buf


class Comm:
    """orbfe_comm: the library's own communicator for the sharded brute force (NCCL bootstrap + mapped peer buffers)."""

    def __init__(self, unique_id, nranks, rank, device, max_queries):
        uid = np.ascontiguousarray(unique_id, np.uint8)
        assert uid.size == COMM_ID_BYTES
        h = C.c_void_p()
        check(lib().orbfe_comm_create(_p(uid), nranks, rank, device, max_queries, C.byref(h)))
        self._h, self.nranks, self.rank, self.device = h, nranks, rank, device

    def close(self):
        if getattr(self, "_h", None):
            lib().orbfe_comm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self):
        n, r, p, v = _I(), _I(), _I(), _I()
        check(lib().orbfe_comm_info(self._h, C.byref(n), C.byref(r), C.byref(p), C.byref(v)))
        return {"nranks": n.value, "rank": r.value, "peer_memory": bool(p.value), "nccl_version": v.value}

    def status(self):
        check(lib().orbfe_comm_status(self._h))

    def set_timing(self, enable=True):
        check(lib().orbfe_comm_set_timing(self._h, int(enable)))

    def last_times(self):
        a, b = _F(), _F()
        check(lib().orbfe_comm_last_times(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def hamming_bruteforce_sharded(self, matcher, d_query, nq, d_map_shard, nm_local, index_base, d_best, d_second,
                                   exchange=EXCHANGE_AUTO):
        """Device pointers (ints); asynchronous on the matcher's stream; collective over the ranks."""
        check(lib().orbfe_hamming_bruteforce_sharded(matcher._h, self._h, d_query, nq, d_map_shard, nm_local,
                                                     index_base, d_best, d_second, exchange))


class Vocabulary:
    """DBoW2 vocabulary tree on the device (orbfe_vocabulary_*): ORBVocabulary::loadFromTextFile / transform."""

    def __init__(self, handle, device):
        self._h, self.device = handle, device
        info = np.zeros(6, np.int32)
        check(lib().orbfe_vocabulary_info(self._h, _p(info)))
        self.k, self.L, self.scoring, self.weighting, self.n_nodes, self.n_words = [int(x) for x in info]

    @classmethod
    def from_arrays(cls, k, L, scoring, weighting, parent, is_leaf, desc, weight, device=0):
        parent = np.ascontiguousarray(parent, np.int32)
        is_leaf = np.ascontiguousarray(is_leaf, np.uint8)
        desc = _u8(desc).reshape(-1, 32)
        weight = np.ascontiguousarray(weight, np.float64)
        h = C.c_void_p()
        check(lib().orbfe_vocabulary_create(device, k, L, scoring, weighting, len(parent), _p(parent), _p(is_leaf),
                                            _p(desc), _p(weight), C.byref(h)))
        return cls(h, device)

    @classmethod
    def load_text(cls, path, device=0):
        h = C.c_void_p()
        check(lib().orbfe_vocabulary_load_text(device, os.fsencode(path), C.byref(h)))
        return cls(h, device)

    def close(self):
        if getattr(self, "_h", None):
            lib().orbfe_vocabulary_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def transform(self, desc, levelsup, device_ptr=None, n=None):
        """Returns (bow_ids, bow_vals, fv_nodes, fv_offsets, fv_features, word_of, node_of, weight_of); with
        device_ptr (an int address of n x 32 device bytes) the per-feature arrays are None."""
        if device_ptr is None:
            desc = _u8(desc).reshape(-1, 32)
            n = len(desc)
        cap = max(n, 1)
        bow_ids, bow_vals = np.empty(cap, np.uint32), np.empty(cap, np.float64)
        fv_nodes, fv_off, fv_feat = np.empty(cap, np.uint32), np.zeros(cap + 1, np.int32), np.empty(cap, np.int32)
        nb, nf = C.c_int32(), C.c_int32()
        if device_ptr is None:
            word, node, wt = np.empty(cap, np.uint32), np.empty(cap, np.uint32), np.empty(cap, np.float64)
            check(lib().orbfe_vocabulary_transform(self._h, _p(desc), n, levelsup, _p(bow_ids), _p(bow_vals), cap,
                                                   C.byref(nb), _p(fv_nodes), _p(fv_off), _p(fv_feat), cap, C.byref(nf),
                                                   _p(word), _p(node), _p(wt)))
            per = (word[:n], node[:n], wt[:n])
        else:
            check(lib().orbfe_vocabulary_transform_device(self._h, device_ptr, n, levelsup, _p(bow_ids), _p(bow_vals),
                                                          cap, C.byref(nb), _p(fv_nodes), _p(fv_off), _p(fv_feat), cap,
                                                          C.byref(nf)))
            per = (None, None, None)
        nb, nf = nb.value, nf.value
        return (bow_ids[:nb], bow_vals[:nb], fv_nodes[:nf], fv_off[:nf + 1], fv_feat[:fv_off[nf]]) + per
