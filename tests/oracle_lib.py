"""ctypes binding of oracle/libygz_oracle.so (CPU restatement; test infrastructure only)."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(ROOT, "oracle", "libygz_oracle.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])


class Params(C.Structure):
    _fields_ = [("nfeatures", C.c_int), ("scaleFactor", C.c_float), ("nlevels", C.c_int),
                ("iniThFAST", C.c_int), ("minThFAST", C.c_int)]


def params(nfeatures=1000, scale=1.2, levels=8, ini=20, mn=7):
    return Params(nfeatures, scale, levels, ini, mn)


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(LIB_PATH)
        _lib.ygzo_cossin_check_libm.restype = C.c_long
        _lib.ygzo_cossin_check_libm.argtypes = [C.c_float, C.c_float, C.c_long]
        _lib.ygzo_round_half_even.argtypes = [C.c_float]
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _img(a):
    a = np.ascontiguousarray(a, np.uint8)
    return a, a.shape[1], a.shape[0]


def tables(p):
    L = p.nlevels
    sc, isc, s2, is2 = (np.zeros(L, np.float32) for _ in range(4))
    npl = np.zeros(L, np.int32)
    um = np.zeros(16, np.int32)
    lib().ygzo_tables(C.byref(p), _p(sc), _p(isc), _p(s2), _p(is2), _p(npl), _p(um))
    return {"scale": sc, "inv_scale": isc, "sigma2": s2, "inv_sigma2": is2, "n_per_level": npl, "umax": um}


def level_sizes(p, w, h):
    lw = np.zeros(p.nlevels, np.int32)
    lh = np.zeros(p.nlevels, np.int32)
    lib().ygzo_level_sizes(C.byref(p), w, h, _p(lw), _p(lh))
    return [(int(a), int(b)) for a, b in zip(lw, lh)]


def resize(src, dw, dh):
    src, w, h = _img(src)
    dst = np.zeros((dh, dw), np.uint8)
    lib().ygzo_resize(_p(src), w, h, w, _p(dst), dw, dh)
    return dst


def blur7(src):
    src, w, h = _img(src)
    dst = np.zeros_like(src)
    lib().ygzo_blur7(_p(src), w, h, w, _p(dst))
    return dst


def border101(src, b):
    src, w, h = _img(src)
    dst = np.zeros((h + 2 * b, w + 2 * b), np.uint8)
    lib().ygzo_border101(_p(src), w, h, w, b, _p(dst))
    return dst


def fast(img, threshold, nms=True):
    img, w, h = _img(img)
    cap = w * h
    xs, ys, rs = (np.zeros(cap, np.int32) for _ in range(3))
    n = lib().ygzo_fast(_p(img), w, h, w, threshold, int(nms), cap, _p(xs), _p(ys), _p(rs))
    return xs[:n].copy(), ys[:n].copy(), rs[:n].copy()


def fast_atan2(y, x):
    y = np.ascontiguousarray(y, np.float32)
    x = np.ascontiguousarray(x, np.float32)
    out = np.zeros_like(y)
    lib().ygzo_fast_atan2(_p(y), _p(x), y.size, _p(out))
    return out


def cossin(a):
    a = np.ascontiguousarray(a, np.float32)
    c = np.zeros_like(a)
    s = np.zeros_like(a)
    lib().ygzo_cossin(_p(a), a.size, _p(c), _p(s))
    return c, s


def pyramid(p, img):
    img, w, h = _img(img)
    sizes = level_sizes(p, w, h)
    out = np.zeros(sum(a * b for a, b in sizes), np.uint8)
    lib().ygzo_pyramid(C.byref(p), _p(img), w, h, w, _p(out))
    lv, off = [], 0
    for lw, lh in sizes:
        lv.append(out[off:off + lw * lh].reshape(lh, lw))
        off += lw * lh
    return lv


def fast_candidates(lvl, ini=20, mn=7):
    lvl, w, h = _img(lvl)
    cap = w * h // 4 + 16
    xs = np.zeros(cap, np.int16)
    ys = np.zeros(cap, np.int16)
    rs = np.zeros(cap, np.uint8)
    n = lib().ygzo_fast_candidates(_p(lvl), w, h, w, ini, mn, cap, _p(xs), _p(ys), _p(rs))
    return xs[:n].copy(), ys[:n].copy(), rs[:n].copy()


def distribute(xs, ys, rs, min_x, max_x, min_y, max_y, N):
    xs = np.ascontiguousarray(xs, np.int16)
    ys = np.ascontiguousarray(ys, np.int16)
    rs = np.ascontiguousarray(rs, np.uint8)
    cap = len(xs) + 8
    out = np.zeros(cap, np.int32)
    n = lib().ygzo_distribute(_p(xs), _p(ys), _p(rs), len(xs), min_x, max_x, min_y, max_y, N, cap, _p(out))
    return out[:n].copy()


def ic_angles(lvl, xs, ys):
    lvl, w, h = _img(lvl)
    xs = np.ascontiguousarray(xs, np.int32)
    ys = np.ascontiguousarray(ys, np.int32)
    out = np.zeros(len(xs), np.float32)
    lib().ygzo_ic_angles(_p(lvl), w, h, w, _p(xs), _p(ys), len(xs), _p(out))
    return out


def descriptors(blurred, xs, ys, angles):
    blurred, w, h = _img(blurred)
    xs = np.ascontiguousarray(xs, np.int32)
    ys = np.ascontiguousarray(ys, np.int32)
    angles = np.ascontiguousarray(angles, np.float32)
    out = np.zeros((len(xs), 32), np.uint8)
    lib().ygzo_descriptors(_p(blurred), w, h, w, _p(xs), _p(ys), _p(angles), len(xs), _p(out))
    return out


def extract(p, img, cap=None):
    img, w, h = _img(img)
    cap = cap or (p.nfeatures + 64) * 2
    kps = np.zeros(cap, KP_DTYPE)
    desc = np.zeros((cap, 32), np.uint8)
    n = lib().ygzo_extract(C.byref(p), _p(img), w, h, w, cap, _p(kps), _p(desc))
    assert n <= cap
    return kps[:n].copy(), desc[:n].copy()


def extract_batch(p, frames, threads=1, cap=None, want_output=True):
    frames = np.ascontiguousarray(frames, np.uint8)
    n, h, w = frames.shape
    cap = cap or (p.nfeatures + 64)
    counts = np.zeros(n, np.int32)
    kps = np.zeros((n, cap), KP_DTYPE) if want_output else None
    desc = np.zeros((n, cap, 32), np.uint8) if want_output else None
    lib().ygzo_extract_batch(C.byref(p), _p(frames), n, w, h, cap, threads, _p(counts),
                             _p(kps) if want_output else None, _p(desc) if want_output else None)
    return counts, kps, desc
