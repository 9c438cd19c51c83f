"""The DSO-style keypoint variant (ORBextractor::operator()(Frame*, ..., DSO_KEYPOINT),
src/ORBextractor.cc:1031-1127 + 1275-1386) on the GPU against the reference's own code (oracle/_ref/ygz_ref_orb --dso)
and the committed golden vectors it produced (tests/golden/dso_*.npz, tools/make_golden_dso.py)."""
import os
import zlib

import numpy as np
import pytest

import refcli
import synth
from ygz import Y

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def existing_keypoints(ref_kps, n, seed, w, h):
    """Keypoints as direct tracking leaves them in frame->mvKeys: sub-pixel level-0 positions on several octaves."""
    rng = np.random.default_rng(seed)
    e = np.zeros(n, Y.KP_DTYPE)
    e["x"] = rng.uniform(40, w - 40, n).astype(np.float32)
    e["y"] = rng.uniform(40, h - 40, n).astype(np.float32)
    e["octave"] = rng.integers(0, 4, n)
    e["size"] = 7.0; e["angle"] = -1.0; e["class_id"] = -1
    if ref_kps is not None and len(ref_kps):     # half of them sit exactly on corners the detector would pick: occupancy
        k = min(n // 2, len(ref_kps))
        pick = rng.choice(len(ref_kps), k, replace=False)
        e["x"][:k] = ref_kps["x"][pick]; e["y"][:k] = ref_kps["y"][pick]; e["octave"][:k] = 0
    return e


def same(got, exp):
    assert len(got[0]) == len(exp["kps"])
    assert got[0].tobytes() == exp["kps"].tobytes()
    assert np.array_equal(got[1], exp["desc"])


@pytest.mark.skipif(not refcli.available(), reason="oracle/_ref/ygz_ref_orb not built")
@pytest.mark.parametrize("w,h,nf,kind,seed", [(752, 480, 1000, "textured", 0), (640, 480, 1000, "mixed", 1),
                                              (752, 480, 300, "textured", 2), (1241, 376, 2000, "textured", 3),
                                              (640, 480, 1500, "lowcontrast", 4)])
def test_dso_stream_vs_reference(w, h, nf, kind, seed):
    """One camera stream: mnGridSize carries from frame to frame (18 -> scalar FAST-10, then >= 22 -> the SSE2 variant)."""
    mk = {"textured": synth.textured_frame, "mixed": synth.mixed_frame,
          "lowcontrast": lambda a, b, s: synth.textured_frame(a, b, s, contrast=0.15)}[kind]
    frames = np.stack([mk(w, h, seed * 10 + k) for k in range(4)])
    first = refcli.run_ref_dso(frames, nfeatures=nf)
    exist = [existing_keypoints(first[k]["kps"], 150 + 40 * k, seed * 7 + k, w, h) for k in range(4)]
    exp = refcli.run_ref_dso(frames, exist, nfeatures=nf)
    ex = Y.Extractor(nf)
    g = -1
    for k in range(4):
        got, ga = ex.extract_dso(frames[k], [exist[k]], [g])
        same(got[0], exp[k])
        assert ga[0] == exp[k]["grid_after"]
        g = ga[0]
        assert len(got[0][0]) > len(exist[k]) + 50
    ex.close()


@pytest.mark.skipif(not refcli.available(), reason="oracle/_ref/ygz_ref_orb not built")
def test_dso_batch_of_independent_frames():
    w, h, nf = 752, 480, 1000
    frames = np.stack([synth.textured_frame(w, h, 40 + k) for k in range(5)])
    grids = [-1, 23, 33, 48, 12]
    exp = [refcli.run_ref_dso(frames[k], None, nfeatures=nf, grid=grids[k])[0] for k in range(5)]
    ex = Y.Extractor(nf, max_batch=8)
    got, ga = ex.extract_dso(frames, None, grids)
    for k in range(5):
        same(got[k], exp[k])
        assert ga[k] == exp[k]["grid_after"]
    ex.close()


@pytest.mark.parametrize("name", ["dso_euroc752", "dso_mixed640"])
def test_dso_golden(name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    mk = synth.textured_frame if int(g["textured"]) else synth.mixed_frame
    frames = np.stack([mk(int(g["w"]), int(g["h"]), int(g["seed"]) + k) for k in range(3)])
    assert [zlib.crc32(f.tobytes()) for f in frames] == g["frame_crc32"].tolist()
    ex = Y.Extractor(int(g["nfeatures"]))
    grid = -1
    for k in range(len(frames)):
        e = g["exist_%d" % k].view(Y.KP_DTYPE).reshape(-1)
        got, ga = ex.extract_dso(frames[k], [e], [grid])
        assert got[0][0].tobytes() == g["kps_%d" % k].tobytes()
        assert np.array_equal(got[0][1], g["desc_%d" % k])
        assert ga[0] == int(g["grid_after"][k])
        grid = ga[0]
    ex.close()
