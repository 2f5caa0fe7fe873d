"""The C-ABI library loads on a machine without a GPU and exports every symbol include/orbfe.h declares; compute
entry points fail loudly (no CPU fallback) when no device is present."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    so = os.path.join(ROOT, "my_orb_slam2_b200", "liborbfe.so")
    if not os.path.exists(so):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "my_orb_slam2_b200", "csrc"), "-j8"])
    from my_orb_slam2_b200 import orbfe
    return orbfe.lib()


def _declared():
    txt = open(os.path.join(ROOT, "include", "orbfe.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(orbfe_[a-z0-9_]+)\s*\(", txt)))


def test_every_declared_symbol_is_exported_and_bound(lib):
    from my_orb_slam2_b200 import orbfe
    names = _declared()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(orbfe.SYMBOLS) == names  # the ctypes table covers exactly the header


def test_struct_layouts_match_header(lib):
    from my_orb_slam2_b200 import orbfe
    assert orbfe.KEYPOINT_DTYPE.itemsize == 28  # cv::KeyPoint
    assert ctypes.sizeof(orbfe.ExtractorParams) == 20
    assert ctypes.sizeof(orbfe.ExtractorInfo) == 4 + 4 + 4 * 64 + 4 * 16 + 4 * 16 + 4
    assert b"sm_100a" in lib.orbfe_version()


def test_no_cpu_fallback(lib):
    from my_orb_slam2_b200 import orbfe
    if orbfe.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(orbfe.OrbfeError):
        orbfe.Extractor(1000, 1.2, 8, 20, 7, 640, 480)
    with pytest.raises(orbfe.OrbfeError):
        orbfe.Matcher()


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "my_orb_slam2_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cc")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "pyoracle" not in txt and "orb_oracle" not in txt and "oracle/" not in txt, f


def test_compat_matcher_compiles_without_gpu():
    """compat/ORBmatcher.cc (the reference's signatures over the C ABI) compiles and links against liborbfe.so."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    subprocess.check_call(["make", "-C", os.path.join(root, "compat")])
    assert os.path.exists(os.path.join(root, "compat", "_build", "test_compat_matcher"))
