"""Pins oracle/minicv (the OpenCV primitives the reference calls) bit-for-bit to cv2 4.13.0.

Reference call sites: src/ORBextractor.cc:1139 (resize), :1141-1146 (copyMakeBorder), :765-768
(FAST), :100 (fastAtan2), :1010 (GaussianBlur), :80,111,115-116 (cvRound).
"""
import numpy as np
import pytest

import oracle_lib as O
import synth

cv2 = pytest.importorskip("cv2")


def _rand_img(w, h, seed, smooth=False):
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 256, size=(h, w), dtype=np.uint8)
    if smooth:
        a = cv2.GaussianBlur(a, (5, 5), 1.2)
    return a


@pytest.mark.parametrize("w,h", [(640, 480), (752, 480), (1241, 376), (1280, 720), (97, 61)])
def test_resize_chain_matches_cv2(w, h):
    p = O.params()
    img = synth.textured_frame(w, h, 1) if w > 100 else _rand_img(w, h, 2)
    sizes = O.level_sizes(p, w, h)
    prev = img
    for lw, lh in sizes[1:]:
        if lw < 2 or lh < 2:
            break
        got = O.resize(prev, lw, lh)
        ref = cv2.resize(prev, (lw, lh), interpolation=cv2.INTER_LINEAR)
        assert np.array_equal(got, ref)
        prev = got


@pytest.mark.parametrize("sw,sh,dw,dh", [(752, 480, 376, 240), (376, 240, 188, 120), (100, 80, 50, 40),
                                         (333, 222, 200, 111), (64, 48, 77, 59), (50, 40, 50, 40), (31, 17, 30, 16)])
def test_resize_misc_scales_incl_exact_2x(sw, sh, dw, dh):
    img = _rand_img(sw, sh, sw + dh)
    assert np.array_equal(O.resize(img, dw, dh), cv2.resize(img, (dw, dh), interpolation=cv2.INTER_LINEAR))


@pytest.mark.parametrize("w,h", [(752, 480), (131, 97), (16, 9), (7, 7), (5, 3), (640, 1)])
def test_blur7_matches_cv2(w, h):
    for seed, smooth in ((0, False), (1, True)):
        img = _rand_img(w, h, seed, smooth)
        ref = cv2.GaussianBlur(img, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
        assert np.array_equal(O.blur7(img), ref)
    for v in (0, 255):
        img = np.full((h, w), v, np.uint8)
        assert np.array_equal(O.blur7(img), cv2.GaussianBlur(img, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101))


def test_border101_matches_cv2():
    img = _rand_img(53, 31, 4)
    assert np.array_equal(O.border101(img, 19), cv2.copyMakeBorder(img, 19, 19, 19, 19, cv2.BORDER_REFLECT_101))


@pytest.mark.parametrize("th", [7, 20, 0, 100])
@pytest.mark.parametrize("shape", [(160, 120), (42, 42), (36, 41), (7, 7), (9, 30)])
def test_fast_matches_cv2(th, shape):
    w, h = shape
    for seed, smooth in ((0, False), (1, True), (2, True)):
        img = _rand_img(w, h, seed + 10 * th, smooth)
        for nms in (True, False):
            det = cv2.FastFeatureDetector_create(threshold=th, nonmaxSuppression=nms,
                                                 type=cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
            kps = det.detect(img)
            xs, ys, rs = O.fast(img, th, nms)
            assert len(kps) == len(xs)
            if nms:  # without NMS OpenCV reports response 0-or-score depending on path; positions only
                ref = [(int(k.pt[0]), int(k.pt[1]), int(k.response)) for k in kps]
                assert ref == list(zip(xs.tolist(), ys.tolist(), rs.tolist()))
            else:
                assert [(int(k.pt[0]), int(k.pt[1])) for k in kps] == list(zip(xs.tolist(), ys.tolist()))


def test_fast_on_textured_level_and_subview():
    img = synth.textured_frame(320, 240, 3)
    for th in (20, 7):
        det = cv2.FastFeatureDetector_create(threshold=th, nonmaxSuppression=True)
        sub = img[16:58, 46:88]
        kps = det.detect(np.ascontiguousarray(sub))
        xs, ys, rs = O.fast(sub, th, True)
        assert [(int(k.pt[0]), int(k.pt[1]), int(k.response)) for k in kps] == list(zip(xs.tolist(), ys.tolist(), rs.tolist()))
        assert len(xs) > 0


def test_fast_atan2_matches_cv2():
    rng = np.random.default_rng(0)
    m = rng.integers(-2 ** 22, 2 ** 22, size=(200000, 2)).astype(np.float32)
    m[:100] = 0
    m[100:200, 0] = 0
    m[200:300, 1] = 0
    m[300:400, 1] = m[300:400, 0]
    m[400:500, 1] = -m[400:500, 0]
    got = O.fast_atan2(m[:, 0], m[:, 1])
    # the reference calls the SCALAR cv::fastAtan2 (src/ORBextractor.cc:100); cv2.phase (vectorised,
    # FMA-contracted in this build) is a different code path and is deliberately not the yardstick
    ref = np.array([cv2.fastAtan2(float(y), float(x)) for y, x in m], np.float32)
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))


def test_cvround_half_even():
    for v, r in ((0.5, 0), (1.5, 2), (2.5, 2), (-0.5, 0), (-1.5, -2), (3.4999, 3), (-2.5, -2)):
        assert O.lib().ygzo_round_half_even(v) == r


def test_cosf_sinf_restatement_equals_libm_sampled():
    # Full-range check (every float in [0, 6.3]) is test_cosf_sinf_exhaustive (slow).
    assert O.lib().ygzo_cossin_check_libm(0.0, 6.3, 257) == 0
    assert O.lib().ygzo_cossin_check_libm(0.78, 0.79, 1) == 0
    assert O.lib().ygzo_cossin_check_libm(1.5, 1.6, 1) == 0


def test_cosf_sinf_exhaustive():
    # every float in [0, 2*pi]: ~1.08e9 values, ~13 s on one core
    assert O.lib().ygzo_cossin_check_libm(0.0, 6.2832, 1) == 0


def undistort_maps(w, h, k1=-0.28, k2=0.07):
    """cv::initUndistortRectifyMap(K, D, Mat(), K, size, CV_16SC2, map1, map2) as Frame::ComputeImagePyramid calls it
    (src/Frame.cc:778-783), for a EuRoC-like camera scaled to the image."""
    K = np.array([[0.61 * w, 0, 0.49 * w], [0, 0.95 * h, 0.52 * h], [0, 0, 1]], np.float64)
    D = np.array([k1, k2, 0.0002, 0.00002, 0.0])
    return cv2.initUndistortRectifyMap(K, D, None, K, (w, h), cv2.CV_16SC2)


@pytest.mark.parametrize("w,h", [(752, 480), (641, 479), (200, 150)])
def test_remap_matches_cv2(w, h):
    img = synth.textured_frame(w, h, 3)
    m1, m2 = undistort_maps(w, h)
    assert np.array_equal(O.remap(img, m1, m2), cv2.remap(img, m1, m2, cv2.INTER_LINEAR))
    rng = np.random.default_rng(1)           # arbitrary maps, taps off every edge
    r1 = rng.integers(-6, max(w, h) + 6, (h, w, 2)).astype(np.int16)
    r2 = rng.integers(0, 1024, (h, w)).astype(np.uint16)
    assert np.array_equal(O.remap(img, r1, r2), cv2.remap(img, r1, r2, cv2.INTER_LINEAR))


def test_mat_float_expressions_vs_cv2():
    """oracle/minicv/minicv_matops.hpp (the cv::Mat expressions of the Sim3 overloads of src/ORBmatcher.cc, used when the
    reference's own file is compiled into oracle/_ref/libygz_ref_match.so) against the installed OpenCV: gemm with and without the
    transposition flag and alpha, norm, dot; scaling = element * (float)alpha as cv::Mat::convertTo does."""
    import ref_match_lib as R
    if not R.available():
        pytest.skip("oracle/_ref/libygz_ref_match.so needs the reference tree at build time")
    import ctypes as C
    rng = np.random.default_rng(3)
    for k in range(300):
        Rm = rng.normal(0, 1 + 10 * (k % 3), (3, 3)).astype(np.float32)
        x = rng.normal(0, 5, (3, 1)).astype(np.float32); t = rng.normal(0, 2, (3, 1)).astype(np.float32)
        s = np.float32(rng.uniform(0.3, 3.0))
        out = np.zeros(38, np.float32)
        R.lib().ygzr_matops_probe(R._p(Rm), R._p(x), R._p(t), C.c_float(s), R._p(out))
        assert out[0:3].tobytes() == cv2.gemm(Rm, x, 1.0, t, 1.0).astype(np.float32).tobytes()
        assert out[3:6].tobytes() == cv2.gemm(Rm, t, -1.0, None, 0.0, flags=cv2.GEMM_1_T).astype(np.float32).tobytes()
        assert out[26:29].tobytes() == cv2.gemm(Rm, t, -1.0, None, 0.0).astype(np.float32).tobytes()
        assert out[6] == np.float32(cv2.norm(x)) and out[7] == np.float32(float(np.dot(x[:, 0].astype(np.float64), t[:, 0].astype(np.float64))))
        assert out[8:17].tobytes() == (Rm.T * np.float32(1.0 / float(s))).astype(np.float32).tobytes()
        assert out[17:26].tobytes() == (Rm * s).astype(np.float32).tobytes()
        assert out[29:38].tobytes() == (Rm * np.float32(1.0 / float(s))).astype(np.float32).tobytes()
