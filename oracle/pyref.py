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
