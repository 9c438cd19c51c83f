"""Pins the CPU oracle (oracle/orb_oracle.cpp) to the reference's own src/ORBextractor.cc:
(a) against committed golden vectors produced by oracle/_ref/ygz_ref_orb (tools/make_golden.py);
(b) live against that binary, when present, on more seeds / shapes / parameters."""
import zlib

import numpy as np
import pytest

import golden_cases
import oracle_lib as O
import refcli
import synth


@pytest.mark.parametrize("name", golden_cases.names())
def test_oracle_matches_golden(name):
    g = golden_cases.load(name)
    p = O.params(g["nfeatures"], g["scale"], g["levels"], g["ini"], g["min"])
    kps, desc = O.extract(p, g["img"])
    assert len(kps) == len(g["kps"])
    assert kps.tobytes() == g["kps"].tobytes()          # positions, size, angle (bit-exact), response, octave
    assert np.array_equal(desc, g["desc"])
    pyr = O.pyramid(p, g["img"])
    assert [zlib.crc32(l.tobytes()) for l in pyr] == [int(c) for c in g["pyr_crc32"]]
    assert [[l.shape[1], l.shape[0]] for l in pyr] == g["pyr_sizes"].tolist()


def test_golden_has_expected_structure():
    g = golden_cases.load("tum640")
    assert np.bincount(g["kps"]["octave"]).tolist() == [217, 181, 151, 127, 106, 88, 73, 61]
    g = golden_cases.load("euroc752_yaml")   # level 3 is 94x60: nRows == 0 -> no cells (Examples/Monocular/EuRoC.yaml:35-38)
    assert np.bincount(g["kps"]["octave"], minlength=4).tolist()[3] == 0


def test_tables():
    t = O.tables(O.params(1000))
    assert t["n_per_level"].tolist() == [217, 181, 151, 126, 105, 87, 73, 60]
    assert O.tables(O.params(2000))["n_per_level"].tolist() == [434, 362, 302, 251, 209, 175, 145, 122]
    assert t["umax"].tolist() == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
    assert O.level_sizes(O.params(), 752, 480) == [(752, 480), (627, 400), (522, 333), (435, 278), (363, 231),
                                                    (302, 193), (252, 161), (210, 134)]


needs_ref = pytest.mark.skipif(not refcli.available(), reason="oracle/_ref/ygz_ref_orb not built (needs /root/reference)")


@needs_ref
@pytest.mark.parametrize("w,h,nf,sc,lv,ini,mn,seed,kind", [
    (640, 480, 1000, 1.2, 8, 20, 7, 100, "t"), (752, 480, 1000, 1.2, 8, 20, 7, 101, "t"),
    (752, 480, 500, 1.2, 8, 20, 7, 102, "m"), (1241, 376, 2000, 1.2, 8, 20, 7, 103, "t"),
    (400, 300, 1500, 1.5, 5, 30, 10, 104, "t"), (320, 480, 800, 1.2, 8, 20, 7, 105, "m"),
    (640, 480, 1000, 1.2, 8, 12, 3, 106, "l"), (333, 211, 5000, 1.3, 6, 20, 7, 107, "t"),
])
def test_oracle_matches_reference_live(w, h, nf, sc, lv, ini, mn, seed, kind):
    img = {"t": synth.textured_frame(w, h, seed), "m": synth.mixed_frame(w, h, seed),
           "l": synth.textured_frame(w, h, seed, contrast=0.15)}[kind]
    outs, _ = refcli.run_ref(img, nfeatures=nf, scale=sc, levels=lv, ini=ini, mn=mn, pyr=True)
    p = O.params(nf, sc, lv, ini, mn)
    kps, desc = O.extract(p, img)
    assert kps.tobytes() == outs[0]["kps"].tobytes()
    assert np.array_equal(desc, outs[0]["desc"])
    for a, b in zip(O.pyramid(p, img), outs[0]["pyr"]):
        assert np.array_equal(a, b)


@needs_ref
def test_reference_frame_overload_equals_inputarray_overload():
    # src/ORBextractor.cc:1031-1127 on a fresh Frame (N == 0) == :970-1028
    img = synth.textured_frame(640, 480, 55)
    a, _ = refcli.run_ref(img)
    b, _ = refcli.run_ref(img, frame_api=True)
    assert a[0]["kps"].tobytes() == b[0]["kps"].tobytes() and np.array_equal(a[0]["desc"], b[0]["desc"])


@needs_ref
def test_reference_multithreaded_batch_is_deterministic():
    frames = synth.frame_batch(320, 240, 6, seed0=40, distinct=3)
    a, _ = refcli.run_ref(frames, nfeatures=400, threads=1)
    b, _ = refcli.run_ref(frames, nfeatures=400, threads=4)
    for x, y in zip(a, b):
        assert x["kps"].tobytes() == y["kps"].tobytes() and np.array_equal(x["desc"], y["desc"])
    counts, kps, desc = O.extract_batch(O.params(400), frames, threads=3, cap=520)
    for f in range(len(frames)):
        assert kps[f, :counts[f]].tobytes() == a[f]["kps"].tobytes()
        assert np.array_equal(desc[f, :counts[f]], a[f]["desc"])


@needs_ref
@pytest.mark.parametrize("name", ["dso_euroc752", "dso_mixed640"])
def test_reference_dso_mode_reproduces_golden(name):
    """The DSO_KEYPOINT goldens (tools/make_golden_dso.py) are what the reference binary produces today."""
    g = np.load(golden_cases.GOLDEN_DIR + "/" + name + ".npz")
    mk = synth.textured_frame if int(g["textured"]) else synth.mixed_frame
    frames = np.stack([mk(int(g["w"]), int(g["h"]), int(g["seed"]) + k) for k in range(3)])
    assert [zlib.crc32(f.tobytes()) for f in frames] == g["frame_crc32"].tolist()
    exist = [g["exist_%d" % k].view(refcli.KP_DTYPE).reshape(-1) for k in range(3)]
    outs = refcli.run_ref_dso(frames, exist, nfeatures=int(g["nfeatures"]))
    for k in range(3):
        assert outs[k]["kps"].tobytes() == g["kps_%d" % k].tobytes() and np.array_equal(outs[k]["desc"], g["desc_%d" % k])
        assert outs[k]["grid_after"] == int(g["grid_after"][k])
        new = outs[k]["kps"][len(exist[k]):]
        assert (new["octave"] == 0).all() and (new["size"] == 7).all() and len(new) > 500
