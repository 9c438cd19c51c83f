"""GPU parity of every ORBmatcher search, Frame::ComputeStereoMatches, MapPoint::ComputeDistinctiveDescriptors and
FindDirectProjection + Align2D against golden vectors made by the REFERENCE'S OWN code (src/ORBmatcher.cc, src/Align.cc and the
hot-path bodies of Frame / KeyFrame / MapPoint compiled unmodified: tests/golden/match_golden.npz, tools/make_golden_match.py),
through the C ABI.  Integer results equal; floats equal by their bytes.  Where the reference library travelled to the box
(oracle/_ref/libygz_ref_match.so) it is also run live on the same inputs."""
import numpy as np
import pytest

import match_cases as MC
import ref_match_lib as R
from test_ref_match import GOLDEN, KEYS, golden_case
from ygz import Y

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def matcher():
    m = Y.Matcher()
    yield m
    m.close()


@pytest.fixture(scope="module")
def golden():
    return np.load(GOLDEN)


@pytest.mark.parametrize("key", KEYS)
def test_kernels_equal_reference_golden(matcher, golden, key):
    d = MC.build_inputs(MC.BY_KEY[key])
    want, der = golden_case(golden, key)
    got = MC.run_gpu(Y, matcher, MC.apply_derived(d, der))
    assert MC.same([np.asarray(x) for x in got], want), key
    if R.available():
        live, _ = MC.run_ref(MC.build_inputs(MC.BY_KEY[key]))
        assert MC.same([np.asarray(x) for x in live], want), key + " (live reference)"
