"""ctypes binding of libygzorb_cuda.so (include/ygzorb.h) for the tests, bench.py and smoke().

Harness only: the product is the shared library and the C++ host classes under host/.  Loading
fails loudly when the library is missing — there is no Python or CPU fallback.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libygzorb_cuda.so")

IN_DEVICE, OUT_DEVICE, KEEP_STAGES, NO_SYNC = 1, 2, 4, 8

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])


class Params(C.Structure):
    _fields_ = [("nfeatures", C.c_int), ("scaleFactor", C.c_float), ("nlevels", C.c_int),
                ("iniThFAST", C.c_int), ("minThFAST", C.c_int)]


class YgzError(RuntimeError):
    pass


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise YgzError("libygzorb_cuda.so is not built (run `make -C orb-ygz-slam_b200` or __graft_entry__.build())")
        L = C.CDLL(LIB_PATH)
        L.ygzorb_last_error.restype = C.c_char_p
        L.ygzorb_version.restype = C.c_char_p
        L.ygzorb_extractor_stream.restype = C.c_void_p
        L.ygzorb_extractor_stream.argtypes = [C.c_void_p]
        L.ygzorb_extractor_launch_count.restype = C.c_longlong
        L.ygzorb_extractor_launch_count.argtypes = [C.c_void_p]
        L.ygzorb_extractor_destroy.restype = None
        L.ygzorb_extractor_destroy.argtypes = [C.c_void_p]
        L.ygzorb_extract_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_size_t,
                                           C.c_uint, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ygzorb_compute_pyramid_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                                   C.c_size_t, C.c_uint]
        _lib = L
    return _lib


def check(rc):
    if rc < 0:
        raise YgzError("ygzorb error %d: %s" % (rc, lib().ygzorb_last_error().decode()))
    return rc


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def device_count():
    return check(lib().ygzorb_device_count())


class Extractor:
    """Mirror of ygz::ORBextractor (include/ORBextractor.h:53-57) over the C ABI."""

    def __init__(self, nfeatures=1000, scale=1.2, levels=8, ini=20, mn=7, max_batch=1, device=0):
        self.params = Params(nfeatures, scale, levels, ini, mn)
        self.max_batch = max_batch
        h = C.c_void_p()
        check(lib().ygzorb_extractor_create(C.byref(self.params), 0, 0, max_batch, device, C.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            lib().ygzorb_extractor_destroy(self.h)
            self.h = None

    __del__ = close

    @property
    def nlevels(self):
        return self.params.nlevels

    def tables(self):
        L = self.nlevels
        sc, isc, s2, is2 = (np.zeros(L, np.float32) for _ in range(4))
        npl = np.zeros(L, np.int32)
        um = np.zeros(16, np.int32)
        check(lib().ygzorb_extractor_tables(self.h, _p(sc), _p(isc), _p(s2), _p(is2), _p(npl), _p(um)))
        return {"scale": sc, "inv_scale": isc, "sigma2": s2, "inv_sigma2": is2, "n_per_level": npl, "umax": um}

    def level_sizes(self, w, h):
        lw = np.zeros(self.nlevels, np.int32)
        lh = np.zeros(self.nlevels, np.int32)
        check(lib().ygzorb_extractor_level_sizes(self.h, w, h, _p(lw), _p(lh)))
        return [(int(a), int(b)) for a, b in zip(lw, lh)]

    def max_keypoints(self, w, h):
        return check(lib().ygzorb_extractor_max_keypoints(self.h, w, h))

    def extract_batch(self, frames, flags=0):
        """frames: uint8 (n, h, w) host array.  Returns list of (kps, desc) per frame."""
        frames = np.ascontiguousarray(frames, np.uint8)
        if frames.ndim == 2:
            frames = frames[None]
        n, h, w = frames.shape
        cap = self.max_keypoints(w, h)
        counts = np.zeros(max(n, 1), np.int32)
        kps = np.zeros((max(n, 1), cap), KP_DTYPE)
        desc = np.zeros((max(n, 1), cap, 32), np.uint8)
        check(lib().ygzorb_extract_batch(self.h, _p(frames), n, w, h, w, w * h, flags, cap, _p(counts), _p(kps), _p(desc)))
        return [(kps[f, :counts[f]].copy(), desc[f, :counts[f]].copy()) for f in range(n)]

    def extract(self, img, flags=0):
        return self.extract_batch(img, flags)[0]

    def extract_raw(self, ptr, n, w, h, stride, frame_stride, flags, cap, counts_ptr, kps_ptr, desc_ptr):
        """Pointer-level call (device or pinned buffers owned by the caller)."""
        check(lib().ygzorb_extract_batch(self.h, C.c_void_p(ptr), n, w, h, stride, frame_stride, flags, cap,
                                         C.c_void_p(counts_ptr), C.c_void_p(kps_ptr), C.c_void_p(desc_ptr)))

    def compute_pyramid(self, frames, flags=0):
        frames = np.ascontiguousarray(frames, np.uint8)
        if frames.ndim == 2:
            frames = frames[None]
        n, h, w = frames.shape
        self._keep = frames
        check(lib().ygzorb_compute_pyramid_batch(self.h, _p(frames), n, w, h, w, w * h, flags))
        self._wh = (w, h)

    def pyramid(self, frame, w, h):
        sizes = self.level_sizes(w, h)
        out = np.zeros(sum(a * b for a, b in sizes), np.uint8)
        check(lib().ygzorb_get_pyramid(self.h, frame, _p(out)))
        lv, off = [], 0
        for lw, lh in sizes:
            lv.append(out[off:off + lw * lh].reshape(lh, lw))
            off += lw * lh
        return lv

    def pyramid_level(self, frame, level, w, h):
        lw, lh = self.level_sizes(w, h)[level]
        out = np.zeros((lh, lw), np.uint8)
        check(lib().ygzorb_get_pyramid_level(self.h, frame, level, _p(out)))
        return out

    def blurred_level(self, frame, level, w, h):
        lw, lh = self.level_sizes(w, h)[level]
        out = np.zeros((lh, lw), np.uint8)
        check(lib().ygzorb_get_blurred_level(self.h, frame, level, _p(out)))
        return out

    def candidates(self, frame, level, w, h):
        lw, lh = self.level_sizes(w, h)[level]
        cap = lw * lh // 4 + 64
        xs = np.zeros(cap, np.int16)
        ys = np.zeros(cap, np.int16)
        rs = np.zeros(cap, np.uint8)
        key = np.zeros(cap, np.uint32)
        n = check(lib().ygzorb_get_candidates(self.h, frame, level, cap, _p(xs), _p(ys), _p(rs), _p(key)))
        order = np.argsort(key[:n], kind="stable")
        return xs[:n][order], ys[:n][order], rs[:n][order]

    def sync(self):
        check(lib().ygzorb_extractor_sync(self.h))

    def stream(self):
        return lib().ygzorb_extractor_stream(self.h)

    def launch_count(self):
        return int(lib().ygzorb_extractor_launch_count(self.h))

    def set_profiling(self, on):
        check(lib().ygzorb_extractor_set_profiling(self.h, int(on)))

    def stage_ms(self):
        ms = np.zeros(8, np.float32)
        check(lib().ygzorb_extractor_stage_ms(self.h, _p(ms), 8))
        return ms


def resize_linear(src, dw, dh, device=0):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.zeros((dh, dw), np.uint8)
    check(lib().ygzorb_resize_linear_u8(_p(src), src.shape[1], src.shape[0], _p(dst), dw, dh, device))
    return dst


def gaussian_blur7(src, device=0):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.zeros_like(src)
    check(lib().ygzorb_gaussian_blur7_u8(_p(src), src.shape[1], src.shape[0], _p(dst), device))
    return dst
