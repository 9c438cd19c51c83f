"""GPU parity of SURVEY.md 8(f) row f4 through the C ABI against the CPU oracle (oracle/direct_oracle.cpp)."""
import numpy as np
import pytest

import oracle_lib as O
from test_direct_oracle import synth_observations
from ygz import Y

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def matcher():
    m = Y.Matcher()
    yield m
    m.close()


@pytest.mark.parametrize("n_points,seed,big", [(1, 0, None), (7, 1, None), (500, 2, None), (40, 3, 700), (3000, 4, None)])
def test_distinctive_descriptors_vs_oracle(matcher, n_points, seed, big):
    """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:211-271): chosen row, its median and mDescriptor, bit-exact;
    points without observations keep their descriptor; `big` exceeds the kernel's shared-memory staging (512 rows)."""
    desc, start = synth_observations(n_points, seed, big=big)
    marker = np.full((n_points, 32), 0x5A, np.uint8)
    e = O.distinctive_descriptors(desc, start, marker.copy())
    g = matcher.distinctive_descriptors(desc, start, marker.copy())
    for a, b in zip(g, e):
        assert np.array_equal(a, b)


def test_distinctive_descriptors_empty(matcher):
    idx, med, out = matcher.distinctive_descriptors(np.zeros((0, 32), np.uint8), np.zeros(4, np.int32))
    assert list(idx) == [-1, -1, -1] and np.all(med == np.iinfo(np.int32).max)


import synth  # noqa: E402


def _gpu_case(c):
    pyrs = [Y.Pyramid(k["levels"]) for k in c["kfs"]]
    cur = Y.Pyramid(c["cur_levels"])
    kfs = [dict(k, pyramid=p) for k, p in zip(c["kfs"], pyrs)]
    return kfs, cur, pyrs + [cur]


@pytest.mark.parametrize("seed,n", [(0, 1), (1, 33), (2, 400), (3, 3000)])
def test_find_direct_projection_vs_oracle(matcher, seed, n):
    """ORBmatcher::FindDirectProjection (src/ORBmatcher.cc:1573-1602) + ygz::Align2D (src/Align.cc:8-105): search level and
    success flag equal, aligned pixel positions bit-exact floats (same evaluation order, no contraction); tolerance 0."""
    c = synth.direct_projection_case(seed, n=n)
    e_px, e_lv, e_ok = O.find_direct_projection(**c)
    kfs, cur, keep = _gpu_case(c)
    args = {k: v for k, v in c.items() if k not in ("kfs", "cur_levels")}
    g_px, g_lv, g_ok = matcher.find_direct_projection(kfs, cur, **args)
    for p in keep:
        p.close()
    assert np.array_equal(g_lv, e_lv) and np.array_equal(g_ok, e_ok)
    assert g_px.tobytes() == e_px.tobytes()
    if n >= 400:
        assert 0.2 < e_ok.mean() < 0.95 and (e_lv > 0).any()


def test_pyramid_snapshot_of_extractor():
    """ygzorb_pyramid_from_extractor: the device-to-device snapshot holds the extractor's pyramid (== the oracle's), continuous."""
    img = synth.textured_frame(752, 480, 5)
    ex = Y.Extractor(1000)
    ex.extract(img)
    snap = Y.Pyramid(extractor=ex, frame=0)
    other = synth.textured_frame(752, 480, 6)
    ex.extract(other)                     # the snapshot must not follow the extractor's next frame
    ref = synth.pyramid_of(img)
    assert snap.sizes() == [(a.shape[1], a.shape[0]) for a in ref]
    for l, a in enumerate(ref):
        assert np.array_equal(snap.level(l), a)
    snap.close(); ex.close()


def _align_handles(matcher, c):
    ref, cur = Y.Pyramid(c["ref_levels"]), Y.Pyramid(c["cur_levels"])
    sa = Y.SparseAlign(matcher)
    sa.begin(ref, cur, c["kp_xy"], c["valid"], c["world_pos"], c["ref_pose"], c["focal"], c["cur_intr"], c["inv_scale"])
    return sa, ref, cur


@pytest.mark.parametrize("level,n", [(1, 700), (3, 2500), (7, 64)])
def test_sparse_align_hot_loops_vs_oracle(matcher, level, n):
    """precomputeReferencePatches + computeResiduals (src/SparseImageAlign.cc:57-225): the caches are bit-exact (per-feature
    float arithmetic in the reference's order); H / Jres / chi2 are sums over features, reduced on the device by a fixed tree in
    double: tolerance 1e-6 relative against the oracle's double accumulation of the same float partials, 2e-4 against the
    reference's sequential float sums.  n_meas exact."""
    c = synth.sparse_align_case(2, n=n)
    args = dict(ref_levels=c["ref_levels"], cur_levels=c["cur_levels"], level=level, scale=float(c["inv_scale"][level]), focal=c["focal"],
                cur_intr=c["cur_intr"], ref_pose=c["ref_pose"], t_cur_ref=c["t_cur_ref"], kp_xy=c["kp_xy"], valid=c["valid"],
                world_pos=c["world_pos"])
    eH, eJ, echi2, enm, evis, epatch, ejac = O.sparse_align_residuals(accumulate=1, **args)
    fH, fJ, fchi2, _, _, _, _ = O.sparse_align_residuals(accumulate=0, **args)
    sa, ref, cur = _align_handles(matcher, c)
    sa.precompute(level)
    vis, patch, jac = sa.caches()
    H, J, chi2, nm = sa.residuals(level, c["t_cur_ref"])
    sa.close(); ref.close(); cur.close()
    assert np.array_equal(vis, evis) and patch.tobytes() == epatch.tobytes() and jac.tobytes() == ejac.tobytes()
    assert nm == enm and nm > 0
    assert np.abs(H - eH).max() <= 1e-6 * np.abs(eH).max() and np.abs(J - eJ).max() <= 1e-6 * np.abs(eJ).max()
    assert abs(chi2 - echi2) <= 1e-6 * echi2
    assert np.abs(H - fH).max() <= 2e-4 * np.abs(fH).max() and np.abs(J - fJ).max() <= 2e-4 * np.abs(fJ).max()
    assert np.allclose(H, H.T)


@pytest.mark.parametrize("seed,n", [(1, 600), (3, 2000)])
def test_sparse_align_run_vs_oracle(matcher, seed, n):
    """SparseImgAlign::run (one kernel launch): TCR within 1e-4 of the oracle's (float Gauss-Newton whose sums differ in the last
    bits), the same number of tracked features, and the known motion of the synthetic case recovered."""
    c = synth.sparse_align_case(seed, n=n)
    args = {k: v for k, v in c.items() if k != "t_true"}
    eT, en, estats = O.sparse_align_run(max_level=7, min_level=1, accumulate=1, **args)
    sa, ref, cur = _align_handles(matcher, c)
    T, nt, stats = sa.run(7, 1, c["t_cur_ref"])
    sa.close(); ref.close(); cur.close()
    assert nt == en and nt > n // 2
    assert np.abs(T - eT).max() < 1e-4
    assert np.abs(T[4:] - c["t_true"]).max() < 2e-3
    assert abs(stats[0] - estats[0]) <= 3


def test_sparse_align_no_features(matcher):
    c = synth.sparse_align_case(0, n=8)
    ref, cur = Y.Pyramid(c["ref_levels"]), Y.Pyramid(c["cur_levels"])
    sa = Y.SparseAlign(matcher)
    sa.begin(ref, cur, np.zeros((0, 2), np.float32), np.zeros(0, np.uint8), np.zeros((0, 3), np.float32), c["ref_pose"], c["focal"],
             c["cur_intr"], c["inv_scale"])
    T, nt, _ = sa.run(7, 1, c["t_cur_ref"])
    assert nt == 0 and np.array_equal(T, c["t_cur_ref"])
    sa.close(); ref.close(); cur.close()
