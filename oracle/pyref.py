"""ctypes binding of oracle/_ref/libref_orbslam2.so: the UNMODIFIED reference sources (/root/reference/src/*.cc)
compiled against the OpenCV stand-in of oracle/ref_shim/ (recipe: oracle/Makefile, driver: oracle/ref_driver.cc).

TEST INFRASTRUCTURE ONLY: used by tests/ (to pin the oracle restatement and, through it, the CUDA path to the
reference's literal code) and by bench.py's cpu_baseline / --impl reference legs (kind "reference").  The product
package never imports this module.  The .so is built in the authoring container (where /root/reference exists) and
travels to the GPU box prebuilt; nothing here reads /root/reference at run time.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from .pyoracle import KEYPOINT_DTYPE

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "libref_orbslam2.so")

MODE_DETERMINISTIC = 0   # creation-order quadtree ties (monotonic node arena) + correctly rounded cosf/sinf/logf
MODE_LIBM = 1            # system libm cosf / sinf / logf, as the reference literally runs
MODE_HEAP_TIES = 2       # plain malloc: quadtree ties by heap address, as the reference literally runs
MODE_AS_IS = MODE_LIBM | MODE_HEAP_TIES

_lib = None


def available():
    return os.path.exists(_SO)


def build():
    """(Re)build when the reference is present; otherwise the prebuilt library must already be there."""
    if os.path.exists("/root/reference/src/ORBextractor.cc"):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not available():
            build()
        if not available():
            raise RuntimeError("oracle/_ref/libref_orbslam2.so is missing and /root/reference is not here to build it")
        L = C.CDLL(_SO)
        P = C.c_void_p
        L.ref_set_mode.argtypes = [C.c_int]
        L.ref_get_mode.restype = C.c_int
        L.ref_list_node_size.restype = C.c_size_t
        L.ref_extractor_create.restype = P
        L.ref_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.ref_extractor_destroy.argtypes = [P]
        L.ref_extract.restype = C.c_int
        L.ref_extract.argtypes = [P, P, C.c_int, C.c_int, C.c_int, P, P, C.c_int]
        L.ref_extractor_level.restype = P
        L.ref_extractor_level.argtypes = [P, C.c_int, P, P, P]
        L.ref_extractor_tables.argtypes = [P, P, P, P, P]
        L.ref_descriptor_distance.restype = C.c_int
        L.ref_descriptor_distance.argtypes = [P, P]
        L.ref_stereo_matches.restype = C.c_int
        L.ref_stereo_matches.argtypes = [P, P, C.c_float, C.c_float, P, P]
        L.ref_extract_pair_threads.argtypes = [P, P, P, P, C.c_int, C.c_int, C.c_int]
        L.ref_extractor_results.restype = C.c_int
        L.ref_extractor_results.argtypes = [P, P, P, C.c_int]
        L.ref_search_by_projection_map.restype = C.c_int
        L.ref_search_by_projection_map.argtypes = [P, P, C.c_float, C.c_float, P]
        L.ref_search_by_projection_last.restype = C.c_int
        L.ref_search_by_projection_last.argtypes = [P, P, P, C.c_float, C.c_int, C.c_int, P]
        L.ref_search_for_triangulation.restype = C.c_int
        L.ref_search_for_triangulation.argtypes = [P, P, P, C.c_float, C.c_float, C.c_int, C.c_int, P, C.c_int]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def set_mode(mode):
    lib().ref_set_mode(int(mode))


class Extractor:
    """ORB_SLAM2::ORBextractor itself (include/ORBextractor.h:45-108, src/ORBextractor.cc), same call shape as
    oracle.pyoracle.Extractor."""

    def __init__(self, nfeatures, scale_factor, nlevels, ini_th, min_th):
        self.nlevels = nlevels
        self.nfeatures = nfeatures
        self._h = lib().ref_extractor_create(nfeatures, scale_factor, nlevels, ini_th, min_th)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().ref_extractor_destroy(self._h)
            self._h = None

    def extract(self, image):
        image = np.ascontiguousarray(image, np.uint8)
        cap = self.nfeatures + 4 * self.nlevels + 64
        while True:
            kps = np.zeros(cap, KEYPOINT_DTYPE)
            desc = np.zeros((cap, 32), np.uint8)
            n = lib().ref_extract(self._h, _p(image), image.shape[1], image.shape[0], image.shape[1], _p(kps),
                                  _p(desc), cap)
            if n <= cap:
                return kps[:n].copy(), desc[:n].copy()
            cap = n

    def level(self, lv):
        w, h, st = C.c_int(), C.c_int(), C.c_int()
        ptr = lib().ref_extractor_level(self._h, lv, C.byref(w), C.byref(h), C.byref(st))
        buf = (C.c_uint8 * (st.value * h.value)).from_address(ptr)
        return np.frombuffer(buf, np.uint8).reshape(h.value, st.value)[:, :w.value].copy()

    def tables(self):
        out = [np.zeros(self.nlevels, np.float32) for _ in range(4)]
        lib().ref_extractor_tables(self._h, *[_p(a) for a in out])
        return out

    def results(self):
        """Keypoints / descriptors the handle holds (after extract() or extract_pair_threads())."""
        cap = self.nfeatures + 4 * self.nlevels + 64
        while True:
            kps = np.zeros(cap, KEYPOINT_DTYPE)
            desc = np.zeros((cap, 32), np.uint8)
            n = lib().ref_extractor_results(self._h, _p(kps), _p(desc), cap)
            if n <= cap:
                return kps[:n].copy(), desc[:n].copy()
            cap = n


def descriptor_distance(a, b):
    a = np.ascontiguousarray(a, np.uint8)
    b = np.ascontiguousarray(b, np.uint8)
    return lib().ref_descriptor_distance(_p(a), _p(b))


def extract_pair_threads(left_ex, right_ex, left, right):
    """The extraction step of the stereo Frame constructor: two std::threads (src/Frame.cc:89-92)."""
    left = np.ascontiguousarray(left, np.uint8)
    right = np.ascontiguousarray(right, np.uint8)
    lib().ref_extract_pair_threads(left_ex._h, right_ex._h, _p(left), _p(right), left.shape[1], left.shape[0],
                                   left.shape[1])
    return left_ex.results(), right_ex.results()


def stereo_matches(left_ex, right_ex, mbf, mb):
    """Frame::ComputeStereoMatches on what the two handles hold from their last extraction -> (uRight, depth, n)."""
    n_left = len(left_ex.results()[0])
    ur = np.full(max(n_left, 1), -1, np.float32)
    dp = np.full(max(n_left, 1), -1, np.float32)
    n = lib().ref_stereo_matches(left_ex._h, right_ex._h, np.float32(mbf), np.float32(mb), _p(ur), _p(dp))
    return ur[:n_left], dp[:n_left], n


# The matcher entry points take the SAME view structs as oracle.pyoracle (make_frame_view, ...): pass those objects.
def _ref(v):
    return C.byref(v.struct) if hasattr(v, "struct") else C.byref(v)


def search_by_projection_map(frame, mps, th, nnratio):
    out = np.full(frame.struct.n if hasattr(frame, "struct") else frame.n, -1, np.int32)
    n = lib().ref_search_by_projection_map(_ref(frame), _ref(mps), th, nnratio, _p(out))
    return out, n


def search_by_projection_last(cur, cam, last, th, mono, check_ori):
    out = np.full(cur.struct.n if hasattr(cur, "struct") else cur.n, -1, np.int32)
    n = lib().ref_search_by_projection_last(_ref(cur), _ref(cam), _ref(last), th, int(mono), int(check_ori), _p(out))
    return out, n


def search_for_triangulation(kf1, kf2, F12, ex, ey, only_stereo, check_ori):
    n1 = kf1.struct.n if hasattr(kf1, "struct") else kf1.n
    pairs = np.zeros((max(n1, 1), 2), np.int32)
    F12 = np.ascontiguousarray(F12, np.float32)
    n = lib().ref_search_for_triangulation(_ref(kf1), _ref(kf2), _p(F12), ex, ey, int(only_stereo), int(check_ori),
                                           _p(pairs), n1)
    return pairs[:n].copy()
