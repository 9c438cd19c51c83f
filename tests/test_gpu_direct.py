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
