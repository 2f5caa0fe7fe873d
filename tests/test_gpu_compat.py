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
