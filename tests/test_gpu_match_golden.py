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


EXACT_KEYS = [c[1] for c in MC.CASES if c[0] not in MC.GPU_TOLERANCE_KINDS]


@pytest.mark.parametrize("key", EXACT_KEYS)
def test_kernels_equal_reference_golden(matcher, golden, key):
    d = MC.build_inputs(MC.BY_KEY[key])
    want, der = golden_case(golden, key)
    got = MC.run_gpu(Y, matcher, MC.apply_derived(d, der))
    assert MC.same([np.asarray(x) for x in got], want), key
    if R.available():
        live, _ = MC.run_ref(MC.build_inputs(MC.BY_KEY[key]))
        assert MC.same([np.asarray(x) for x in live], want), key + " (live reference)"


# ---- SparseImgAlign: golden vectors from the reference's own src/SparseImageAlign.cc + include/NLSSolver_impl.hpp ---------------
def _align_handles(matcher, c):
    ref, cur = Y.Pyramid(c["ref_levels"]), Y.Pyramid(c["cur_levels"])
    sa = Y.SparseAlign(matcher)
    sa.begin(ref, cur, c["kp_xy"], c["valid"], c["world_pos"], c["ref_pose"], c["focal"], c["cur_intr"], c["inv_scale"])
    return sa, ref, cur


@pytest.mark.parametrize("key", [c[1] for c in MC.CASES if c[0] == "sares"])
def test_sparse_align_hot_loops_vs_reference_golden(matcher, golden, key):
    """precomputeReferencePatches + computeResiduals (src/SparseImageAlign.cc:57-225) against what the reference's own code left in
    visible_fts_, ref_patch_cache_, jacobian_cache_ (bit-exact: per-feature float arithmetic in the reference's order), n_meas_
    (exact) and H_, Jres_, chi2 (sums over features: sequential float in the reference, a fixed tree in double on the device —
    tolerance 2e-4 relative to the largest entry)."""
    d = MC.build_inputs(MC.BY_KEY[key])
    want, _ = golden_case(golden, key)
    wH, wJ, wchi2, wnm, wvis, wpatch, wjac = want
    c = d["c"]
    sa, ref, cur = _align_handles(matcher, c)
    sa.precompute(d["level"])
    vis, patch, jac = sa.caches()
    H, J, chi2, nm = sa.residuals(d["level"], c["t_cur_ref"])
    sa.close(); ref.close(); cur.close()
    assert np.array_equal(vis, wvis) and patch.tobytes() == wpatch.tobytes() and jac.tobytes() == wjac.tobytes()
    assert nm == int(wnm) and nm > 0
    assert np.abs(H - wH).max() <= 2e-4 * np.abs(wH).max() and np.abs(J - wJ).max() <= 2e-4 * np.abs(wJ).max()
    assert abs(chi2 / nm - float(wchi2)) <= 2e-4 * float(wchi2)


@pytest.mark.parametrize("key", [c[1] for c in MC.CASES if c[0] == "sarun"])
def test_sparse_align_run_vs_reference_golden(matcher, golden, key):
    """SparseImgAlign::run, one kernel launch, against the TCR the reference's own run() returned from the same starting pose:
    quaternion and translation within 1e-4 (ten Gauss-Newton steps per level on float sums that differ in the last bits), the
    same return value n_meas_ / 16, final chi2 within 1e-3 relative."""
    d = MC.build_inputs(MC.BY_KEY[key])
    want, der = golden_case(golden, key)
    d = MC.apply_derived(d, der)
    wT, wret, wchi2 = want
    c = d["c"]
    sa, ref, cur = _align_handles(matcher, c)
    T, nt, stats = sa.run(d["max_level"], d["min_level"], c["t_cur_ref"])
    sa.close(); ref.close(); cur.close()
    assert nt == int(wret)
    if int(wret) == 0:          # nothing visible: the pose comes back unchanged
        assert np.abs(T - wT).max() < 1e-6
        return
    assert np.abs(T - wT).max() < 1e-4
    assert abs(stats[1] - float(wchi2)) <= 1e-3 * float(wchi2)
