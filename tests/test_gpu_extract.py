"""GPU parity of the extractor (libygzorb_cuda.so through its C ABI) against the CPU oracle and
the golden vectors made by the reference's own ORBextractor.cc.  Integer/byte outputs bit-exact;
angles compared bit-exact too (tolerance stated by north_star: 1e-4 deg)."""
import zlib

import numpy as np
import pytest

import golden_cases
import oracle_lib as O
import synth
from ygz import Y

pytestmark = pytest.mark.gpu


def _params(g):
    return dict(nfeatures=g["nfeatures"], scale=g["scale"], levels=g["levels"], ini=g["ini"], mn=g["min"])


@pytest.mark.parametrize("sw,sh,dw,dh", [(640, 480, 533, 400), (752, 480, 627, 400), (1241, 376, 1034, 313),
                                          (752, 480, 376, 240), (101, 77, 84, 64), (64, 64, 127, 131), (37, 5, 31, 4)])
def test_resize_kernel_vs_oracle(sw, sh, dw, dh):
    rng = np.random.default_rng(sw * 7 + dh)
    src = rng.integers(0, 256, size=(sh, sw), dtype=np.uint8)
    assert np.array_equal(Y.resize_linear(src, dw, dh), O.resize(src, dw, dh))


@pytest.mark.parametrize("w,h", [(752, 480), (131, 97), (7, 9), (128, 16), (129, 17), (3, 2), (1000, 33)])
def test_blur_kernel_vs_oracle(w, h):
    rng = np.random.default_rng(w + h)
    src = rng.integers(0, 256, size=(h, w), dtype=np.uint8)
    assert np.array_equal(Y.gaussian_blur7(src), O.blur7(src))


def test_tables_match_oracle():
    for nf, sc, lv in [(1000, 1.2, 8), (2000, 1.2, 8), (1000, 2.0, 4), (1500, 1.5, 5)]:
        ex = Y.Extractor(nf, sc, lv)
        t, o = ex.tables(), O.tables(O.params(nf, sc, lv))
        for k in t:
            assert t[k].tobytes() == o[k].tobytes(), k
        assert ex.level_sizes(752, 480) == O.level_sizes(O.params(nf, sc, lv), 752, 480)
        ex.close()


@pytest.mark.parametrize("name", golden_cases.names())
def test_stages_and_extract_vs_golden(name):
    g = golden_cases.load(name)
    img = g["img"]
    h, w = img.shape
    p = O.params(g["nfeatures"], g["scale"], g["levels"], g["ini"], g["min"])
    ex = Y.Extractor(**_params(g))
    kps, desc = ex.extract(img, Y.KEEP_STAGES)
    # K1 pyramid: every level, every pixel + the reference's CRCs
    pyr = ex.pyramid(0, w, h)
    opyr = O.pyramid(p, img)
    for l, (a, b) in enumerate(zip(pyr, opyr)):
        assert a.shape == b.shape and np.array_equal(a, b), "pyramid level %d" % l
    assert [zlib.crc32(l.tobytes()) for l in pyr] == [int(c) for c in g["pyr_crc32"]]
    # K2 FAST candidates (vToDistributeKeys order) and K5 blur, per level
    for l in range(g["levels"]):
        xs, ys, rs = ex.candidates(0, l, w, h)
        oxs, oys, ors = O.fast_candidates(opyr[l], g["ini"], g["min"])
        assert len(xs) == len(oxs), "candidate count level %d" % l
        assert np.array_equal(xs, oxs) and np.array_equal(ys, oys) and np.array_equal(rs, ors), "candidates level %d" % l
        assert np.array_equal(ex.blurred_level(0, l, w, h), O.blur7(opyr[l])), "blur level %d" % l
    # K3/K4/K6 + assembly: the reference's own output
    assert len(kps) == len(g["kps"])
    for fld in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(kps[fld], g["kps"][fld]), fld
    assert np.max(np.abs(kps["angle"] - g["kps"]["angle"])) <= 1e-4   # north_star tolerance ...
    assert kps["angle"].tobytes() == g["kps"]["angle"].tobytes()       # ... and in fact bit-exact
    assert np.array_equal(desc, g["desc"])
    ex.close()


def test_batch_chunking_and_distinct_frames():
    w, h = 320, 240
    frames = np.stack([synth.textured_frame(w, h, 50 + i) for i in range(5)] + [synth.mixed_frame(w, h, 60)])
    p = O.params(500, 1.2, 8, 20, 7)
    ex = Y.Extractor(500, 1.2, 8, 20, 7, max_batch=4)   # 6 frames -> 2 chunks
    outs = ex.extract_batch(frames)
    for f in range(len(frames)):
        okps, odesc = O.extract(p, frames[f])
        assert outs[f][0].tobytes() == okps.tobytes(), "frame %d" % f
        assert np.array_equal(outs[f][1], odesc)
    ex.close()


@pytest.mark.parametrize("w,h,nf,sc,lv,ini,mn,seed,kind", [
    (400, 300, 1500, 1.5, 5, 30, 10, 104, "t"), (320, 480, 800, 1.2, 8, 20, 7, 105, "m"),
    (640, 480, 1000, 1.2, 8, 12, 3, 106, "l"), (333, 211, 5000, 1.3, 6, 20, 7, 107, "t"),
    (96, 80, 100, 1.2, 8, 20, 7, 108, "t"), (1280, 720, 2000, 1.2, 8, 20, 7, 109, "t"),
])
def test_extract_vs_oracle_shapes(w, h, nf, sc, lv, ini, mn, seed, kind):
    img = {"t": synth.textured_frame(w, h, seed), "m": synth.mixed_frame(w, h, seed),
           "l": synth.textured_frame(w, h, seed, contrast=0.15)}[kind]
    ex = Y.Extractor(nf, sc, lv, ini, mn)
    kps, desc = ex.extract(img)
    okps, odesc = O.extract(O.params(nf, sc, lv, ini, mn), img, cap=2 * nf + 512)
    assert len(kps) == len(okps)
    assert kps.tobytes() == okps.tobytes()
    assert np.array_equal(desc, odesc)
    ex.close()


def test_flat_and_empty_images():
    ex = Y.Extractor()
    kps, desc = ex.extract(np.full((480, 640), 90, np.uint8))     # no corners anywhere
    assert len(kps) == 0 and desc.shape == (0, 32)
    counts = np.zeros(1, np.int32)
    rc = Y.lib().ygzorb_extract_batch(ex.h, None, 1, 0, 0, 0, 0, 0, 0, counts.ctypes.data, None, None)
    assert rc == 0 and counts[0] == 0                              # empty image: silent return (:972-973)
    ex.close()


def test_strided_input_and_repeatability():
    w, h = 300, 200
    big = synth.textured_frame(w + 40, h, 77)
    view = np.ascontiguousarray(big[:, 20:20 + w])
    ex = Y.Extractor(400)
    a = ex.extract(view)
    cap = ex.max_keypoints(w, h)
    counts = np.zeros(1, np.int32); kps = np.zeros(cap, Y.KP_DTYPE); desc = np.zeros((cap, 32), np.uint8)
    sub = big[:, 20:]
    Y.check(Y.lib().ygzorb_extract_batch(ex.h, sub.ctypes.data, 1, w, h, big.strides[0], 0, 0, cap, counts.ctypes.data,
                                         kps.ctypes.data, desc.ctypes.data))
    assert counts[0] == len(a[0]) and kps[:counts[0]].tobytes() == a[0].tobytes()
    assert np.array_equal(desc[:counts[0]], a[1])
    b = ex.extract(view)
    assert a[0].tobytes() == b[0].tobytes() and np.array_equal(a[1], b[1])
    ex.close()


@pytest.mark.parametrize("w,h,n", [(752, 480, 24), (750, 300, 40), (640, 480, 20)])
def test_large_batch_per_frame_results(w, h, n):
    """Larger batches (the kernels' batch-size dependent choices: strip resize with several sub-tiles per CTA, 384-thread quadtree)
    give per-frame results equal to the oracle's; 750 px wide frames have 2-byte aligned rows (generic resize and blur paths).
    Textured, mixed-contrast (minThFAST retry cells) and flat frames in one batch."""
    frames = np.stack([synth.textured_frame(w, h, 300 + i) if i % 4 else synth.mixed_frame(w, h, 300 + i) for i in range(n)])
    frames[3] = 90   # a flat frame: every cell retries at minThFAST and finds nothing
    p = O.params(1000, 1.2, 8, 20, 7)
    ex = Y.Extractor(1000, 1.2, 8, 20, 7, max_batch=n)
    outs = ex.extract_batch(frames)
    for f in range(n):
        okps, odesc = O.extract(p, frames[f])
        assert outs[f][0].tobytes() == okps.tobytes(), "frame %d" % f
        assert np.array_equal(outs[f][1], odesc)
    ex.close()


@pytest.mark.parametrize("pipe,max_batch", [(2, 16), (3, 4), (0, 16), (128, 16)])
def test_pipelined_host_chunks(pipe, max_batch):
    """Host-input calls split into pipelined chunks (upload / kernels / download on three streams, two staging
    slots each way) must give exactly the per-frame results, also over back-to-back NO_SYNC calls that reuse
    the slots while earlier work is still in flight."""
    w, h = 320, 240
    frames = np.stack([synth.textured_frame(w, h, 150 + i) if i % 3 else synth.mixed_frame(w, h, 150 + i) for i in range(11)])
    p = O.params(500, 1.2, 8, 20, 7)
    want = [O.extract(p, f) for f in frames]
    ex = Y.Extractor(500, 1.2, 8, 20, 7, max_batch=max_batch)
    ex.set_pipeline(pipe)
    outs = ex.extract_batch(frames)
    for f in range(len(frames)):
        assert outs[f][0].tobytes() == want[f][0].tobytes(), "frame %d" % f
        assert np.array_equal(outs[f][1], want[f][1])
    # three un-synchronised calls into separate host buffers, then one sync
    cap = ex.max_keypoints(w, h)
    bufs = []
    for r in range(3):
        sel = np.ascontiguousarray(frames[r:r + 7])
        counts = np.zeros(7, np.int32); kps = np.zeros((7, cap), Y.KP_DTYPE); desc = np.zeros((7, cap, 32), np.uint8)
        ex.extract_raw(sel.ctypes.data, 7, w, h, w, w * h, Y.NO_SYNC, cap, counts.ctypes.data, kps.ctypes.data, desc.ctypes.data)
        bufs.append((sel, counts, kps, desc))
    ex.sync()
    for r, (_, counts, kps, desc) in enumerate(bufs):
        for i in range(7):
            assert kps[i, :counts[i]].tobytes() == want[r + i][0].tobytes(), (r, i)
            assert np.array_equal(desc[i, :counts[i]], want[r + i][1])
    ex.close()


def test_keep_stages_covers_whole_call():
    """YGZORB_KEEP_STAGES turns the upload pipelining off: the stage taps then hold every frame of the call."""
    w, h = 200, 160
    frames = np.stack([synth.textured_frame(w, h, 170 + i) for i in range(5)])
    ex = Y.Extractor(300, max_batch=8)
    ex.set_pipeline(2)
    ex.extract_batch(frames, Y.KEEP_STAGES)
    for f in range(5):
        assert np.array_equal(ex.pyramid(f, w, h)[0], frames[f])
    ex.close()


@pytest.mark.parametrize("px,py", [(2, 4), (2, 2), (3, 2)])
def test_dot_grid_maxima_density(px, py):
    """Isolated bright dots on a regular grid: every dot is a FAST corner and a strict local maximum, so up to 17
    of 32 consecutive raster pixels survive the NMS (px=2) and a 2x2 block always holds one (px=py=2) — the densest
    inputs for the per-warp lists of k_fast and for the per-level candidate capacity."""
    w, h = 320, 240
    img = np.full((h, w), 40, np.uint8)
    img[::py, ::px] = 220
    ex = Y.Extractor(1000, 1.2, 8, 20, 7)
    kps, desc = ex.extract(img)
    okps, odesc = O.extract(O.params(1000, 1.2, 8, 20, 7), img, cap=4096)
    assert len(okps) > 200
    assert kps.tobytes() == okps.tobytes() and np.array_equal(desc, odesc)
    ex.close()


@pytest.mark.parametrize("w,h", [(752, 480), (641, 479)])
def test_undistort_in_front_of_the_pyramid(w, h):
    """Frame::ComputeImagePyramid's cv::remap (src/Frame.cc:775-790) on the device: level 0 == the oracle's remap (itself
    pinned to cv2.remap), and extraction / DSO extraction of the distorted frame == extraction of the undistorted one."""
    frames = np.stack([synth.textured_frame(w, h, 60 + k) for k in range(3)])
    m1, m2 = synth.undistort_maps(w, h, 0.25, 0.05)
    want = [O.remap(f, m1, m2) for f in frames]
    assert (want[0][0, :] == 0).any() and not np.array_equal(want[0], frames[0])      # the maps do reach outside the image
    ex = Y.Extractor(1000, max_batch=4)
    ex.set_undistort_maps(m1, m2)
    ex.compute_pyramid(frames)
    for k in range(3):
        assert np.array_equal(ex.pyramid(k, w, h)[0], want[k])
    got = ex.extract_batch(frames)
    plain = Y.Extractor(1000, max_batch=4)
    exp = plain.extract_batch(np.stack(want))
    for g, e in zip(got, exp):
        assert g[0].tobytes() == e[0].tobytes() and np.array_equal(g[1], e[1]) and len(g[0]) > 500
    gd, _ = ex.extract_dso(frames[:2])
    ed, _ = plain.extract_dso(np.stack(want[:2]))
    for g, e in zip(gd, ed):
        assert g[0].tobytes() == e[0].tobytes() and np.array_equal(g[1], e[1]) and len(g[0]) > 500
    ex.set_undistort_maps(None, None)                       # and off again
    back = ex.extract_batch(frames)
    ref = plain.extract_batch(frames)
    assert back[0][0].tobytes() == ref[0][0].tobytes()
    ex.close(); plain.close()
