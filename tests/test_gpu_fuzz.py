"""Randomised parity: image size, extractor parameters, image kind (exact ties, saturated pixels, empty cells, corners
in the border band ...) and stereo baseline drawn per seed (tests/fuzz_parity.py); keypoints, descriptors, mvuRight
and mvDepth of the CUDA path must equal the oracle's bit for bit.  ORBFE_FUZZ_N / ORBFE_FUZZ_FIRST widen the sweep
(the sweeps that were run on a B200 are recorded in profiles/)."""
import os

import pytest

import fuzz_parity

pytestmark = pytest.mark.gpu


def test_random_cases_match_oracle(oracle):
    from my_orb_slam2_b200 import frontend, orbfe
    first = int(os.environ.get("ORBFE_FUZZ_FIRST", "0"))
    count = int(os.environ.get("ORBFE_FUZZ_N", "40"))
    ok = refused = 0
    bad = []
    for seed in range(first, first + count):
        d, st = fuzz_parity.run_case(seed, oracle, frontend, orbfe)
        if st == "ok":
            ok += 1
        elif st.startswith("refused"):
            refused += 1
        else:
            bad.append("%s: %s" % (d, st))
    assert not bad, "\n".join(bad)
    assert ok >= count * 0.7, "%d of %d cases refused" % (refused, count)


def test_random_matcher_cases_match_oracle(oracle):
    """Every ORBmatcher member once per seed with drawn thresholds / ratios / windows / motion (tests/fuzz_matchers.py)."""
    import fuzz_matchers
    from my_orb_slam2_b200 import frontend
    first = int(os.environ.get("ORBFE_FUZZ_FIRST", "0"))
    count = int(os.environ.get("ORBFE_FUZZ_MATCHER_N", "25"))
    bad = []
    for seed in range(first, first + count):
        bad += ["seed %d %s" % (seed, name) for name, st, _ in fuzz_matchers.run_case(seed, oracle, frontend) if st != "ok"]
    assert not bad, "\n".join(bad)
