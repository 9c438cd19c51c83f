"""Pins the matcher / direct-tracking restatement (oracle/match_oracle.cpp, oracle/direct_oracle.cpp) to the REFERENCE'S OWN
code: /root/reference/src/ORBmatcher.cc and src/Align.cc compiled whole and unmodified, with the hot-path member functions
of Frame / KeyFrame / MapPoint copied out of src/*.cc at build time (oracle/_ref/libygz_ref_match.so; oracle/Makefile target
refmatch, oracle/ref_match_api.cpp).  Two anchors per case of tests/match_cases.py:
  * the committed golden vectors that library produced (tests/golden/match_golden.npz, tools/make_golden_match.py) — always;
  * the library itself, live — where it exists (the authoring container and the GPU box; skipped elsewhere).
Integer results equal; floats (stereo uRight / depth, vbPrevMatched, aligned pixels) equal by their bytes."""
import os
import sys

import numpy as np
import pytest

import match_cases as MC
import ref_match_lib as R

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
GOLDEN = os.path.join(ROOT, "tests", "golden", "match_golden.npz")
KEYS = [c[1] for c in MC.CASES]


@pytest.fixture(scope="module")
def golden():
    return np.load(GOLDEN)


def golden_case(g, key):
    outs, i = [], 0
    while "%s#o%d" % (key, i) in g.files:
        outs.append(g["%s#o%d" % (key, i)]); i += 1
    der = {k.split("#d_")[1]: g[k] for k in g.files if k.startswith(key + "#d_")}
    return outs, der


def test_golden_covers_every_case(golden):
    have = {k.split("#")[0] for k in golden.files}
    assert have == set(KEYS)


@pytest.mark.parametrize("key", KEYS)
def test_restatement_equals_reference_golden(golden, key):
    from make_golden_match import digest
    d = MC.build_inputs(MC.BY_KEY[key])
    assert digest(d) == bytes(golden[key + "#in"]).hex(), "the seeded inputs differ from those the golden was made from"
    want, der = golden_case(golden, key)
    got = MC.run_oracle(MC.apply_derived(d, der))
    assert MC.same([np.asarray(x) for x in got], want)
    if d["kind"] in ("sbp", "last", "kf", "init", "tri", "bow") and "empty" not in key:
        assert int(want[0]) > 10          # the case exercises the search


@pytest.mark.skipif(not R.available(), reason="oracle/_ref/libygz_ref_match.so needs the reference tree at build time")
@pytest.mark.parametrize("key", KEYS)
def test_reference_live_equals_golden(golden, key):
    d = MC.build_inputs(MC.BY_KEY[key])
    outs, der = MC.run_ref(d)
    want, wder = golden_case(golden, key)
    assert MC.same([np.asarray(x) for x in outs], want)
    assert set(der) == set(wder) and all(np.asarray(der[k]).tobytes() == wder[k].tobytes() for k in der)


@pytest.mark.skipif(not R.available(), reason="oracle/_ref/libygz_ref_match.so needs the reference tree at build time")
def test_descriptor_distance_and_predict_scale_live():
    """DescriptorDistance on random rows; PredictScale over a sweep of distance ratios straddling every level boundary: the
    reference's log(float) (glibc logf under `using namespace std`) and the restatement's choice give the same level."""
    import oracle_lib as O
    rng = np.random.default_rng(0)
    a = rng.integers(0, 256, (200, 32), dtype=np.uint8); b = rng.integers(0, 256, (200, 32), dtype=np.uint8)
    for i in range(200):
        assert R.descriptor_distance(a[i], b[i]) == O.descriptor_distance(a[i], b[i]) == int(np.unpackbits(a[i] ^ b[i]).sum())


@pytest.mark.skipif(not R.available(), reason="oracle/_ref/libygz_ref_match.so needs the reference tree at build time")
def test_reference_batch_leg_equals_restatement():
    """bench.py's CPU leg (the reference's SearchByProjection over a batch, threads over frames) returns what the restatement's
    batch function returns, whatever the thread count and the number of passes."""
    import oracle_lib as O
    import synth
    img, kps, desc = MC.feats()
    sf = MC.tables()[0]
    F, M, n = 5, 700, len(kps)
    cap = n + 8
    K = np.zeros((F, cap), kps.dtype); D = np.zeros((F, cap, 32), np.uint8)
    mp = {k: [] for k in ("proj_x", "proj_y", "proj_xr", "view_cos", "level", "in_view", "has_obs", "desc")}
    for f in range(F):
        K[f, :n] = kps; D[f, :n] = desc
        m, _, _ = synth.projection_case(kps, desc, M, 100 + f)
        for k in mp:
            mp[k].append(m[k])
    mp = {k: np.ascontiguousarray(np.stack(v)) for k, v in mp.items()}
    mp["n_mp"] = np.full(F, M, np.int32)
    counts = np.full(F, n, np.int32)
    want, wnm = O.search_by_projection_batch(counts, mp["n_mp"], K, D, MC.BOUNDS, sf, mp, 1.0, False, 0.8, threads=2)
    for threads, repeat in [(1, 1), (3, 2)]:
        got, nm, sec = R.search_by_projection_batch(counts, mp["n_mp"], K, D, MC.BOUNDS, sf, mp, 1.0, False, 0.8, threads=threads, repeat=repeat)
        assert np.array_equal(got[:, :n], want[:, :n]) and np.array_equal(nm, wnm) and sec > 0
