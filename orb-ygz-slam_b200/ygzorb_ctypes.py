"""ctypes binding of libygzorb_cuda.so (include/ygzorb.h) for the tests, bench.py and smoke().

Harness only: the product is the shared library and the C++ host classes under host/.  Loading
fails loudly when the library is missing — there is no Python or CPU fallback.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("YGZORB_LIB") or os.path.join(HERE, "libygzorb_cuda.so")   # YGZORB_LIB: kernel experiments only

IN_DEVICE, OUT_DEVICE, KEEP_STAGES, NO_SYNC = 1, 2, 4, 8

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])


class Params(C.Structure):
    _fields_ = [("nfeatures", C.c_int), ("scaleFactor", C.c_float), ("nlevels", C.c_int),
                ("iniThFAST", C.c_int), ("minThFAST", C.c_int)]


class YgzError(RuntimeError):
    pass


_lib = None
_live = []   # weak references to every open handle object (Extractor, Matcher, Comm, ...): close_all() for an orderly shutdown


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
        L.ygzorb_extract_dso_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_uint,
                                               C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ygzorb_extractor_set_undistort_maps.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.ygzorb_compute_pyramid_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                                   C.c_size_t, C.c_uint]
        L.ygzorb_hamming_knn2_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_float, C.c_int,
                                                  C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def _track(obj):
    import weakref
    _live.append(weakref.ref(obj))


def close_all():
    """Closes every handle still open, in reverse creation order.  Scripts call it before they return so that handles (and their
    CUDA streams) go before the interpreter tears down whatever else holds device or pinned memory — the order an embedder gets
    when the handles are members destroyed first."""
    for r in reversed(_live):
        o = r()
        if o is not None and getattr(o, "h", None):
            o.close()
    del _live[:]


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
        _track(self)

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

    def extract_dso(self, frames, exist=None, grid=None, cap=None):
        """DSO_KEYPOINT method on independent frames: exist = list of KP_DTYPE arrays (frame->mvKeys on entry) or None,
        grid = list of mnGridSize values before each frame (-1: uninitialised).
        Returns (list of (kps, desc) per frame, list of mnGridSize after each frame)."""
        frames = np.ascontiguousarray(frames, np.uint8)
        if frames.ndim == 2:
            frames = frames[None]
        n, h, w = frames.shape
        exist = exist if exist is not None else [np.zeros(0, KP_DTYPE)] * n
        n_exist = np.array([len(e) for e in exist], np.int32)
        g = np.array(grid if grid is not None else [-1] * n, np.int32)
        cap = cap or int(n_exist.max()) + 3 * (h // 7) * (w // 7)
        counts = np.zeros(n, np.int32)
        kps = np.zeros((n, cap), KP_DTYPE)
        desc = np.zeros((n, cap, 32), np.uint8)
        for i, e in enumerate(exist):
            kps[i, :len(e)] = e
        check(lib().ygzorb_extract_dso_batch(self.h, _p(frames), n, w, h, w, w * h, 0, _p(g), _p(n_exist), cap, _p(counts),
                                             _p(kps), _p(desc)))
        return [(kps[i, :counts[i]].copy(), desc[i, :counts[i]].copy()) for i in range(n)], g.tolist()

    def set_undistort_maps(self, map1, map2):
        """cv::initUndistortRectifyMap(..., CV_16SC2) maps ((h, w, 2) int16, (h, w) uint16); None, None removes them."""
        if map1 is None:
            check(lib().ygzorb_extractor_set_undistort_maps(self.h, None, None, 0, 0))
            return
        m1 = np.ascontiguousarray(map1, np.int16); m2 = np.ascontiguousarray(map2, np.uint16)
        check(lib().ygzorb_extractor_set_undistort_maps(self.h, _p(m1), _p(m2), m2.shape[1], m2.shape[0]))

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

    def set_pipeline(self, frames_per_chunk):
        check(lib().ygzorb_extractor_set_pipeline(self.h, int(frames_per_chunk)))

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


# ---- matcher --------------------------------------------------------------------------------------
class FrameFeatures(C.Structure):
    _fields_ = [("n_frames", C.c_int), ("kp_stride", C.c_int), ("n_kp", C.c_void_p), ("kps", C.c_void_p),
                ("desc", C.c_void_p), ("u_right", C.c_void_p), ("claimed", C.c_void_p),
                ("min_x", C.c_float), ("min_y", C.c_float), ("max_x", C.c_float), ("max_y", C.c_float),
                ("nlevels", C.c_int), ("scale_factors", C.c_void_p)]


class MapPointProj(C.Structure):
    _fields_ = [("mp_stride", C.c_int), ("n_mp", C.c_void_p), ("proj_x", C.c_void_p), ("proj_y", C.c_void_p),
                ("proj_xr", C.c_void_p), ("view_cos", C.c_void_p), ("level", C.c_void_p), ("in_view", C.c_void_p),
                ("has_obs", C.c_void_p), ("desc", C.c_void_p)]


class LastFramePoints(C.Structure):
    _fields_ = [("pt_stride", C.c_int), ("n_pts", C.c_void_p), ("valid", C.c_void_p), ("has_obs", C.c_void_p),
                ("world_pos", C.c_void_p), ("desc", C.c_void_p), ("octave", C.c_void_p), ("angle", C.c_void_p),
                ("pose", C.c_void_p)]


class KeyFramePoints(C.Structure):
    _fields_ = [("pt_stride", C.c_int), ("n_pts", C.c_void_p), ("valid", C.c_void_p), ("world_pos", C.c_void_p),
                ("normal", C.c_void_p), ("desc", C.c_void_p), ("angle", C.c_void_p), ("max_distance", C.c_void_p),
                ("min_distance", C.c_void_p), ("pose", C.c_void_p), ("log_scale_factor", C.c_float),
                ("n_scale_levels", C.c_int)]


class KeyFrameFeatures(C.Structure):
    _fields_ = [("n", C.c_int), ("n_nodes", C.c_int), ("nlevels", C.c_int), ("kps", C.c_void_p), ("desc", C.c_void_p),
                ("u_right", C.c_void_p), ("has_mappoint", C.c_void_p), ("node_ids", C.c_void_p),
                ("node_start", C.c_void_p), ("node_feats", C.c_void_p), ("scale_factors", C.c_void_p),
                ("level_sigma2", C.c_void_p)]


def _ptr(a):
    return a.ctypes.data if a is not None else None


class DirectKeyFrame(C.Structure):
    _fields_ = [("pyramid", C.c_void_p), ("pose_ref", C.c_float * 7), ("t_cur_ref", C.c_float * 7),
                ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float)]


class Pyramid:
    """Device-resident image pyramid (ygzorb_pyramid): from host images or as a snapshot of an extractor's frame."""

    def __init__(self, levels=None, device=0, extractor=None, frame=0):
        L = lib()
        L.ygzorb_pyramid_create.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_uint, C.c_void_p]
        L.ygzorb_pyramid_from_extractor.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.ygzorb_pyramid_destroy.argtypes = [C.c_void_p]
        L.ygzorb_pyramid_destroy.restype = None
        L.ygzorb_pyramid_info.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ygzorb_pyramid_get_level.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        h = C.c_void_p()
        if extractor is not None:
            check(L.ygzorb_pyramid_from_extractor(extractor.h, frame, C.byref(h)))
        else:
            imgs = [np.ascontiguousarray(a, np.uint8) for a in levels]
            ptrs = (C.c_void_p * len(imgs))(*[a.ctypes.data for a in imgs])
            w = np.array([a.shape[1] for a in imgs], np.int32)
            hh = np.array([a.shape[0] for a in imgs], np.int32)
            check(L.ygzorb_pyramid_create(device, ptrs, _ptr(w), _ptr(hh), _ptr(w), len(imgs), 0, C.byref(h)))
        self.h = h
        _track(self)

    def sizes(self):
        n = C.c_int32()
        w, h = np.zeros(16, np.int32), np.zeros(16, np.int32)
        check(lib().ygzorb_pyramid_info(self.h, C.byref(n), _ptr(w), _ptr(h)))
        return [(int(w[l]), int(h[l])) for l in range(n.value)]

    def level(self, l):
        w, h = self.sizes()[l]
        out = np.zeros((h, w), np.uint8)
        check(lib().ygzorb_pyramid_get_level(self.h, l, _ptr(out)))
        return out

    def close(self):
        if self.h:
            lib().ygzorb_pyramid_destroy(self.h)
            self.h = None


class SparseAlign:
    """ygzorb_sparse_align: SparseImgAlign on the device (caches + the fused Gauss-Newton kernel)."""

    def __init__(self, matcher):
        L = lib()
        L.ygzorb_sparse_align_create.argtypes = [C.c_void_p, C.c_void_p]
        L.ygzorb_sparse_align_destroy.argtypes = [C.c_void_p]
        L.ygzorb_sparse_align_destroy.restype = None
        L.ygzorb_sparse_align_begin.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                                C.c_void_p, C.c_float, C.c_void_p, C.c_void_p]
        L.ygzorb_sparse_align_run.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ygzorb_sparse_align_precompute.argtypes = [C.c_void_p, C.c_int]
        L.ygzorb_sparse_align_residuals.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ygzorb_sparse_align_caches.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        h = C.c_void_p()
        check(L.ygzorb_sparse_align_create(matcher.h, C.byref(h)))
        self.h, self.n = h, 0

    def begin(self, ref, cur, kp_xy, valid, world_pos, ref_pose, focal, cur_intr, inv_scale):
        kp_xy = np.ascontiguousarray(kp_xy, np.float32); valid = np.ascontiguousarray(valid, np.uint8)
        world_pos = np.ascontiguousarray(world_pos, np.float32); ref_pose = np.ascontiguousarray(ref_pose, np.float32)
        cur_intr = np.ascontiguousarray(cur_intr, np.float32); inv_scale = np.ascontiguousarray(inv_scale, np.float32)
        self.n = len(valid)
        check(lib().ygzorb_sparse_align_begin(self.h, ref.h, cur.h, self.n, _ptr(kp_xy), _ptr(valid), _ptr(world_pos), _ptr(ref_pose),
                                              C.c_float(focal), _ptr(cur_intr), _ptr(inv_scale)))

    def run(self, max_level, min_level, t_cur_ref):
        T = np.array(t_cur_ref, np.float32).copy()
        n = C.c_int64()
        stats = np.zeros(2, np.float32)
        check(lib().ygzorb_sparse_align_run(self.h, max_level, min_level, _ptr(T), C.byref(n), _ptr(stats)))
        return T, n.value, stats

    def precompute(self, level):
        check(lib().ygzorb_sparse_align_precompute(self.h, level))

    def residuals(self, level, t_cur_ref):
        T = np.ascontiguousarray(t_cur_ref, np.float32)
        H, J, chi2, n = np.zeros((6, 6), np.float32), np.zeros(6, np.float32), C.c_float(), C.c_int64()
        check(lib().ygzorb_sparse_align_residuals(self.h, level, _ptr(T), _ptr(H), _ptr(J), C.byref(chi2), C.byref(n)))
        return H, J, chi2.value, n.value

    def caches(self):
        vis, patch, jac = np.zeros(self.n, np.uint8), np.zeros((self.n, 16), np.float32), np.zeros((self.n, 16, 6), np.float32)
        check(lib().ygzorb_sparse_align_caches(self.h, _ptr(vis), _ptr(patch), _ptr(jac)))
        return vis, patch, jac

    def close(self):
        if self.h:
            lib().ygzorb_sparse_align_destroy(self.h)
            self.h = None


class Matcher:
    """Handle behind ygz::ORBmatcher's Hamming searches and Frame::ComputeStereoMatches."""

    def __init__(self, device=0):
        h = C.c_void_p()
        L = lib()
        L.ygzorb_matcher_destroy.restype = None
        L.ygzorb_matcher_destroy.argtypes = [C.c_void_p]
        L.ygzorb_matcher_stream.restype = C.c_void_p
        L.ygzorb_matcher_stream.argtypes = [C.c_void_p]
        L.ygzorb_matcher_launch_count.restype = C.c_longlong
        L.ygzorb_matcher_launch_count.argtypes = [C.c_void_p]
        L.ygzorb_hamming_knn2.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_float, C.c_int,
                                          C.c_uint, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ygzorb_descriptor_distance.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_uint, C.c_void_p]
        L.ygzorb_distinctive_descriptors.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_uint, C.c_void_p, C.c_void_p,
                                                     C.c_void_p]
        L.ygzorb_search_by_bow.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_uint, C.c_void_p,
                                           C.c_void_p]
        L.ygzorb_bow_transform.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_uint] + \
            [C.c_void_p] * 10
        L.ygzorb_vocabulary_create.argtypes = [C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 5 + [C.c_int, C.c_void_p]
        L.ygzorb_vocabulary_load_binary.argtypes = [C.c_char_p, C.c_int, C.c_void_p]
        L.ygzorb_vocabulary_load_text.argtypes = [C.c_char_p, C.c_int, C.c_void_p]
        L.ygzorb_vocabulary_destroy.argtypes = [C.c_void_p]
        L.ygzorb_vocabulary_destroy.restype = None
        L.ygzorb_vocabulary_info.argtypes = [C.c_void_p, C.c_void_p]
        L.ygzorb_vocabulary_arrays.argtypes = [C.c_void_p] * 6
        L.ygzorb_search_by_projection_last.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_int, C.c_int,
                                                       C.c_uint, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ygzorb_search_by_projection_reloc.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_int, C.c_uint,
                                                        C.c_void_p, C.c_void_p, C.c_void_p]
        L.ygzorb_search_by_projection_sim3.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_uint, C.c_void_p,
                                                       C.c_void_p, C.c_void_p]
        L.ygzorb_search_for_initialization.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_int,
                                                       C.c_uint, C.c_void_p, C.c_void_p]
        L.ygzorb_fuse.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_void_p, C.c_uint, C.c_void_p, C.c_void_p]
        L.ygzorb_search_by_sim3.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_uint,
                                            C.c_void_p, C.c_void_p]
        L.ygzorb_search_by_projection.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_float,
                                                  C.c_uint, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ygzorb_search_for_triangulation.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float,
                                                      C.c_float, C.c_int, C.c_int, C.c_uint, C.c_void_p, C.c_void_p]
        L.ygzorb_stereo_match.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                          C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_uint, C.c_void_p,
                                          C.c_void_p]
        L.ygzorb_measure_popc_peak.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        check(L.ygzorb_matcher_create(device, C.byref(h)))
        self.h = h
        _track(self)

    def close(self):
        if getattr(self, "h", None):
            lib().ygzorb_matcher_destroy(self.h)
            self.h = None

    __del__ = close

    def sync(self):
        check(lib().ygzorb_matcher_sync(self.h))

    def stream(self):
        return lib().ygzorb_matcher_stream(self.h)

    def launch_count(self):
        return int(lib().ygzorb_matcher_launch_count(self.h))

    def set_stream(self, stream):
        L = lib()
        L.ygzorb_matcher_set_stream.argtypes = [C.c_void_p, C.c_void_p]
        check(L.ygzorb_matcher_set_stream(self.h, stream))

    def set_knn_variant(self, v):
        check(lib().ygzorb_matcher_set_knn_variant(self.h, v))

    def popc_peak(self):
        v = C.c_double()
        check(lib().ygzorb_measure_popc_peak(self.h, C.byref(v)))
        return v.value

    def descriptor_distance(self, a, b):
        a = np.ascontiguousarray(a, np.uint8).reshape(-1, 32)
        b = np.ascontiguousarray(b, np.uint8).reshape(-1, 32)
        out = np.zeros(len(a), np.int32)
        check(lib().ygzorb_descriptor_distance(self.h, _ptr(a), _ptr(b), len(a), 0, _ptr(out)))
        return out

    def distinctive_descriptors(self, obs_desc, obs_start, best_desc=None):
        """MapPoint::ComputeDistinctiveDescriptors over a CSR of observation descriptors -> (best_idx, best_median, best_desc)."""
        obs_desc = np.ascontiguousarray(obs_desc, np.uint8).reshape(-1, 32)
        obs_start = np.ascontiguousarray(obs_start, np.int32)
        n = len(obs_start) - 1
        idx, med = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.int32)
        out = np.zeros((max(n, 1), 32), np.uint8) if best_desc is None else np.ascontiguousarray(best_desc, np.uint8)
        check(lib().ygzorb_distinctive_descriptors(self.h, _ptr(obs_desc) if len(obs_desc) else None, _ptr(obs_start), n, 0,
                                                   _ptr(idx), _ptr(med), _ptr(out)))
        return idx[:n], med[:n], out[:n]

    def find_direct_projection(self, kfs, cur, cur_intr, scale, inv_scale, inv_sigma2_1, kf_index, px_ref, octave, world_pos,
                               px_curr):
        """ORBmatcher::FindDirectProjection, batched.  kfs: list of dicts {pyramid: Pyramid, pose_ref[7], t_cur_ref[7], fx, fy,
        cx, cy}; cur: Pyramid of the current frame.  Returns (px_curr_out (n, 2), search_level, success)."""
        L = lib()
        L.ygzorb_find_direct_projection.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                                    C.c_float, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint,
                                                    C.c_void_p, C.c_void_p, C.c_void_p]
        arr = (DirectKeyFrame * max(len(kfs), 1))()
        for a, k in zip(arr, kfs):
            a.pyramid = k["pyramid"].h
            a.pose_ref[:] = [float(x) for x in k["pose_ref"]]
            a.t_cur_ref[:] = [float(x) for x in k["t_cur_ref"]]
            a.fx, a.fy, a.cx, a.cy = (float(k[n]) for n in ("fx", "fy", "cx", "cy"))
        kf_index = np.ascontiguousarray(kf_index, np.int32); octave = np.ascontiguousarray(octave, np.int32)
        px_ref = np.ascontiguousarray(px_ref, np.float32); world_pos = np.ascontiguousarray(world_pos, np.float32)
        out = np.array(px_curr, np.float32).reshape(-1, 2).copy()
        n = len(kf_index)
        intr = np.ascontiguousarray(cur_intr, np.float32)
        scale = np.ascontiguousarray(scale, np.float32); inv_scale = np.ascontiguousarray(inv_scale, np.float32)
        level, ok = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.uint8)
        check(L.ygzorb_find_direct_projection(self.h, arr, len(kfs), cur.h, _ptr(intr), _ptr(scale), _ptr(inv_scale),
                                              C.c_float(inv_sigma2_1), n, _ptr(kf_index), _ptr(px_ref), _ptr(octave),
                                              _ptr(world_pos), 0, _ptr(out), _ptr(level), _ptr(ok)))
        return out, level[:n], ok[:n]

    def hamming_knn2(self, query, train, ratio, th):
        query = np.ascontiguousarray(query, np.uint8).reshape(-1, 32)
        train = np.ascontiguousarray(train, np.uint8).reshape(-1, 32)
        nq = len(query)
        idx, d1, d2, match = (np.zeros(max(nq, 1), np.int32) for _ in range(4))
        check(lib().ygzorb_hamming_knn2(self.h, _ptr(query), nq, _ptr(train) if len(train) else None, len(train), ratio, th,
                                        0, _ptr(idx), _ptr(d1), _ptr(d2), _ptr(match)))
        return idx[:nq], d1[:nq], d2[:nq], match[:nq]

    def hamming_knn2_raw(self, q_ptr, nq, t_ptr, nt, ratio, th, flags, idx_ptr, d1_ptr, d2_ptr, match_ptr):
        check(lib().ygzorb_hamming_knn2(self.h, q_ptr, nq, t_ptr, nt, ratio, th, flags, idx_ptr, d1_ptr, d2_ptr, match_ptr))

    def search_by_projection(self, frames, mps, bounds, scale_factors, th, check_level, nnratio):
        """frames: list of (kps, desc, u_right|None, claimed|None); mps: list of dicts (see tests/synth.projection_case).
        Returns per frame (n_matches, kp_to_mp, mp_best_idx)."""
        nf = len(frames)
        ks = max(max(len(f[0]) for f in frames), 1)
        ms = max(max(len(m["proj_x"]) for m in mps), 1)
        n_kp = np.array([len(f[0]) for f in frames], np.int32)
        n_mp = np.array([len(m["proj_x"]) for m in mps], np.int32)
        kps = np.zeros((nf, ks), KP_DTYPE); desc = np.zeros((nf, ks, 32), np.uint8)
        has_ur = any(f[2] is not None for f in frames); has_cl = any(f[3] is not None for f in frames)
        ur = np.full((nf, ks), -1, np.float32) if has_ur else None
        cl = np.zeros((nf, ks), np.uint8) if has_cl else None
        for i, f in enumerate(frames):
            kps[i, :len(f[0])] = f[0]; desc[i, :len(f[0])] = f[1]
            if f[2] is not None:
                ur[i, :len(f[0])] = f[2]
            if f[3] is not None:
                cl[i, :len(f[0])] = f[3]
        pa = {k: np.zeros((nf, ms), np.float32) for k in ("proj_x", "proj_y", "proj_xr", "view_cos")}
        lv = np.zeros((nf, ms), np.int32); iv = np.zeros((nf, ms), np.uint8); ho = np.zeros((nf, ms), np.uint8)
        md = np.zeros((nf, ms, 32), np.uint8)
        for i, m in enumerate(mps):
            k = len(m["proj_x"])
            for key in pa:
                pa[key][i, :k] = m[key]
            lv[i, :k] = m["level"]; iv[i, :k] = m["in_view"]; ho[i, :k] = m["has_obs"]; md[i, :k] = m["desc"]
        sf = np.ascontiguousarray(scale_factors, np.float32)
        F = FrameFeatures(nf, ks, _ptr(n_kp), _ptr(kps), _ptr(desc), _ptr(ur), _ptr(cl), bounds[0], bounds[1], bounds[2],
                          bounds[3], len(sf), _ptr(sf))
        P = MapPointProj(ms, _ptr(n_mp), _ptr(pa["proj_x"]), _ptr(pa["proj_y"]), _ptr(pa["proj_xr"]), _ptr(pa["view_cos"]),
                         _ptr(lv), _ptr(iv), _ptr(ho), _ptr(md))
        kp_to_mp = np.zeros((nf, ks), np.int32); best = np.zeros((nf, ms), np.int32); nm = np.zeros(nf, np.int32)
        check(lib().ygzorb_search_by_projection(self.h, C.byref(F), C.byref(P), th, int(check_level), nnratio, 0,
                                                _ptr(kp_to_mp), _ptr(best), _ptr(nm)))
        return [(int(nm[i]), kp_to_mp[i, :n_kp[i]].copy(), best[i, :n_mp[i]].copy()) for i in range(nf)]

    def search_by_projection_last(self, frames, lasts, bounds, scale_factors, th, mono, check_level, check_orientation):
        """frames: list of (kps, desc, u_right|None, claimed|None) of the current frames; lasts: list of dicts with
        valid, has_obs, world_pos (n,3), desc, octave, angle, pose (30,) (tests/synth.lastframe_case).
        Returns per frame pair (n_matches, kp_to_pt, pt_best_idx)."""
        nf = len(frames)
        ks = max(max(len(f[0]) for f in frames), 1)
        ps = max(max(len(l["valid"]) for l in lasts), 1)
        n_kp = np.array([len(f[0]) for f in frames], np.int32)
        n_pt = np.array([len(l["valid"]) for l in lasts], np.int32)
        kps = np.zeros((nf, ks), KP_DTYPE); desc = np.zeros((nf, ks, 32), np.uint8)
        has_ur = any(f[2] is not None for f in frames); has_cl = any(f[3] is not None for f in frames)
        ur = np.full((nf, ks), -1, np.float32) if has_ur else None
        cl = np.zeros((nf, ks), np.uint8) if has_cl else None
        for i, f in enumerate(frames):
            kps[i, :len(f[0])] = f[0]; desc[i, :len(f[0])] = f[1]
            if f[2] is not None:
                ur[i, :len(f[0])] = f[2]
            if f[3] is not None:
                cl[i, :len(f[0])] = f[3]
        va = np.zeros((nf, ps), np.uint8); ho = np.zeros((nf, ps), np.uint8); wp = np.zeros((nf, ps, 3), np.float32)
        pd = np.zeros((nf, ps, 32), np.uint8); oc = np.zeros((nf, ps), np.int32); an = np.zeros((nf, ps), np.float32)
        pose = np.zeros((nf, 30), np.float32)
        for i, l in enumerate(lasts):
            k = len(l["valid"])
            va[i, :k] = l["valid"]; ho[i, :k] = l["has_obs"]; wp[i, :k] = l["world_pos"]; pd[i, :k] = l["desc"]
            oc[i, :k] = l["octave"]; an[i, :k] = l["angle"]; pose[i] = l["pose"]
        sf = np.ascontiguousarray(scale_factors, np.float32)
        F = FrameFeatures(nf, ks, _ptr(n_kp), _ptr(kps), _ptr(desc), _ptr(ur), _ptr(cl), bounds[0], bounds[1], bounds[2],
                          bounds[3], len(sf), _ptr(sf))
        L = LastFramePoints(ps, _ptr(n_pt), _ptr(va), _ptr(ho), _ptr(wp), _ptr(pd), _ptr(oc), _ptr(an), _ptr(pose))
        kp_to_pt = np.zeros((nf, ks), np.int32); best = np.zeros((nf, ps), np.int32); nm = np.zeros(nf, np.int32)
        check(lib().ygzorb_search_by_projection_last(self.h, C.byref(F), C.byref(L), th, int(mono), int(check_level),
                                                     int(check_orientation), 0, _ptr(kp_to_pt), _ptr(best), _ptr(nm)))
        return [(int(nm[i]), kp_to_pt[i, :n_kp[i]].copy(), best[i, :n_pt[i]].copy()) for i in range(nf)]

    def _keyframe_views(self, frames, pts, bounds, scale_factors, keep):
        nf = len(frames)
        ks = max(max(len(f[0]) for f in frames), 1)
        ps = max(max(len(l["valid"]) for l in pts), 1)
        n_kp = np.array([len(f[0]) for f in frames], np.int32)
        n_pt = np.array([len(l["valid"]) for l in pts], np.int32)
        kps = np.zeros((nf, ks), KP_DTYPE); desc = np.zeros((nf, ks, 32), np.uint8)
        has_cl = any(len(f) > 2 and f[2] is not None for f in frames)
        has_ur = any(len(f) > 3 and f[3] is not None for f in frames)
        cl = np.zeros((nf, ks), np.uint8) if has_cl else None
        ur = np.full((nf, ks), -1, np.float32) if has_ur else None
        for i, f in enumerate(frames):
            kps[i, :len(f[0])] = f[0]; desc[i, :len(f[0])] = f[1]
            if has_cl and f[2] is not None:
                cl[i, :len(f[0])] = f[2]
            if has_ur and f[3] is not None:
                ur[i, :len(f[0])] = f[3]
        va = np.zeros((nf, ps), np.uint8); wp = np.zeros((nf, ps, 3), np.float32); nr = np.zeros((nf, ps, 3), np.float32)
        pd = np.zeros((nf, ps, 32), np.uint8); an = np.zeros((nf, ps), np.float32)
        mx = np.ones((nf, ps), np.float32); mn = np.ones((nf, ps), np.float32)
        pose = np.zeros((nf, 30), np.float32)
        for i, l in enumerate(pts):
            k = len(l["valid"])
            va[i, :k] = l["valid"]; wp[i, :k] = l["world_pos"]; nr[i, :k] = l["normal"]; pd[i, :k] = l["desc"]
            an[i, :k] = l["angle"]; mx[i, :k] = l["max_distance"]; mn[i, :k] = l["min_distance"]; pose[i] = l["pose"]
        sf = np.ascontiguousarray(scale_factors, np.float32)
        keep.extend([n_kp, n_pt, kps, desc, cl, ur, va, wp, nr, pd, an, mx, mn, pose, sf])
        F = FrameFeatures(nf, ks, _ptr(n_kp), _ptr(kps), _ptr(desc), _ptr(ur), _ptr(cl), bounds[0], bounds[1], bounds[2],
                          bounds[3], len(sf), _ptr(sf))
        L = KeyFramePoints(ps, _ptr(n_pt), _ptr(va), _ptr(wp), _ptr(nr), _ptr(pd), _ptr(an), _ptr(mx), _ptr(mn), _ptr(pose),
                           float(pts[0]["log_scale_factor"]), int(pts[0]["n_scale_levels"]))
        return F, L, n_kp, n_pt, ks, ps

    def search_by_projection_keyframe(self, frames, pts, bounds, scale_factors, th, variant, orb_dist=100, check_orientation=False,
                                      inv_sigma2=None):
        """variant 0: relocalisation SearchByProjection, 1: Sim3 SearchByProjection, 2: Fuse, 3: Fuse with Sim3.
        frames: list of (kps, desc, claimed|None[, u_right|None]); pts: list of dicts with valid, world_pos, normal, desc,
        angle, max_distance, min_distance, pose (30,), log_scale_factor, n_scale_levels (tests/synth.keyframe_points_case).
        Returns per pair (n, kp_to_pt, pt_best_idx) (kp_to_pt is None for Fuse)."""
        keep = []
        variant = int(variant)
        F, L, n_kp, n_pt, ks, ps = self._keyframe_views(frames, pts, bounds, scale_factors, keep)
        nf = len(frames)
        kp_to_pt = np.zeros((nf, ks), np.int32); best = np.zeros((nf, ps), np.int32); nm = np.zeros(nf, np.int32)
        if variant == 1:
            check(lib().ygzorb_search_by_projection_sim3(self.h, C.byref(F), C.byref(L), int(th), 0, _ptr(kp_to_pt), _ptr(best),
                                                         _ptr(nm)))
        elif variant == 0:
            check(lib().ygzorb_search_by_projection_reloc(self.h, C.byref(F), C.byref(L), th, int(orb_dist),
                                                          int(check_orientation), 0, _ptr(kp_to_pt), _ptr(best), _ptr(nm)))
        else:
            is2 = np.ascontiguousarray(inv_sigma2, np.float32) if inv_sigma2 is not None else None
            check(lib().ygzorb_fuse(self.h, C.byref(F), C.byref(L), th, int(variant == 3), _ptr(is2), 0, _ptr(best), _ptr(nm)))
            return [(int(nm[i]), None, best[i, :n_pt[i]].copy()) for i in range(nf)]
        return [(int(nm[i]), kp_to_pt[i, :n_kp[i]].copy(), best[i, :n_pt[i]].copy()) for i in range(nf)]

    def search_for_initialization(self, frames1, frames2, prev_matched, bounds2, window_size, nnratio, check_orientation):
        """frames1 / frames2: lists of (kps, desc); prev_matched: list of (n1, 2) float32 arrays (vbPrevMatched).
        Returns per pair (n_matches, matches12, updated prev_matched)."""
        nf = len(frames1)
        s1 = max(max(len(f[0]) for f in frames1), 1); s2 = max(max(len(f[0]) for f in frames2), 1)
        n1 = np.array([len(f[0]) for f in frames1], np.int32); n2 = np.array([len(f[0]) for f in frames2], np.int32)
        k1 = np.zeros((nf, s1), KP_DTYPE); d1 = np.zeros((nf, s1, 32), np.uint8)
        k2 = np.zeros((nf, s2), KP_DTYPE); d2 = np.zeros((nf, s2, 32), np.uint8)
        pm = np.zeros((nf, s1, 2), np.float32)
        for i in range(nf):
            k1[i, :n1[i]] = frames1[i][0]; d1[i, :n1[i]] = frames1[i][1]
            k2[i, :n2[i]] = frames2[i][0]; d2[i, :n2[i]] = frames2[i][1]
            pm[i, :n1[i]] = prev_matched[i]
        F1 = FrameFeatures(nf, s1, _ptr(n1), _ptr(k1), _ptr(d1), None, None, bounds2[0], bounds2[1], bounds2[2], bounds2[3], 0, None)
        F2 = FrameFeatures(nf, s2, _ptr(n2), _ptr(k2), _ptr(d2), None, None, bounds2[0], bounds2[1], bounds2[2], bounds2[3], 0, None)
        m12 = np.zeros((nf, s1), np.int32); nm = np.zeros(nf, np.int32)
        check(lib().ygzorb_search_for_initialization(self.h, C.byref(F1), C.byref(F2), _ptr(pm), int(window_size), nnratio,
                                                     int(check_orientation), 0, _ptr(m12), _ptr(nm)))
        return [(int(nm[i]), m12[i, :n1[i]].copy(), pm[i, :n1[i]].copy()) for i in range(nf)]

    def search_by_sim3(self, kfs1, kfs2, pts1, pts2, bounds1, bounds2, scale_factors, th):
        """kfs1 / kfs2: lists of (kps, desc); pts1[i] = MapPoints of kfs1[i] (one per keypoint) with the pose that takes them
        into kfs2[i], pts2 the other way round.  Returns per pair (n_found, matches12)."""
        keep = []
        F2, L1, _, n_pt1, _, ps1 = self._keyframe_views(kfs2, pts1, bounds2, scale_factors, keep)
        F1, L2, _, _, _, _ = self._keyframe_views(kfs1, pts2, bounds1, scale_factors, keep)
        nf = len(kfs1)
        m12 = np.zeros((nf, ps1), np.int32); nfound = np.zeros(nf, np.int32)
        check(lib().ygzorb_search_by_sim3(self.h, C.byref(F1), C.byref(F2), C.byref(L1), C.byref(L2), th, 0, _ptr(m12), _ptr(nfound)))
        return [(int(nfound[i]), m12[i, :n_pt1[i]].copy()) for i in range(nf)]

    def search_by_projection_raw(self, F, P, th, check_level, nnratio, flags, kp_to_mp_ptr, best_ptr, nm_ptr):
        check(lib().ygzorb_search_by_projection(self.h, C.byref(F), C.byref(P), th, int(check_level), nnratio, flags,
                                                kp_to_mp_ptr, best_ptr, nm_ptr))

    def search_for_triangulation(self, kf1, kf2, F12, ex, ey, only_stereo, check_orientation):
        keep = []

        def view(kf):
            a = [np.ascontiguousarray(kf["kps"]), np.ascontiguousarray(kf["desc"], np.uint8),
                 np.ascontiguousarray(kf["u_right"], np.float32), np.ascontiguousarray(kf["has_mp"], np.uint8),
                 np.ascontiguousarray(kf["node_ids"], np.uint32), np.ascontiguousarray(kf["node_start"], np.int32),
                 np.ascontiguousarray(kf["node_feats"], np.uint32), np.ascontiguousarray(kf["scale_factors"], np.float32),
                 np.ascontiguousarray(kf["level_sigma2"], np.float32)]
            keep.append(a)
            return KeyFrameFeatures(len(a[0]), len(a[4]), len(a[7]), *[_ptr(x) for x in a])
        v1, v2 = view(kf1), view(kf2)
        F = np.ascontiguousarray(F12, np.float32).reshape(9)
        m12 = np.zeros(max(v1.n, 1), np.int32); nm = np.zeros(1, np.int32)
        check(lib().ygzorb_search_for_triangulation(self.h, C.byref(v1), C.byref(v2), _ptr(F), float(ex), float(ey),
                                                    int(only_stereo), int(check_orientation), 0, _ptr(m12), _ptr(nm)))
        return int(nm[0]), m12[:v1.n].copy()

    def search_by_bow(self, kf1, second, second_is_frame, nnratio, check_orientation):
        """kf dicts: kps, desc, has_mp, node_ids, node_start, node_feats.  Returns (n_matches, matches)."""
        keep = []

        def view(kf):
            a = [np.ascontiguousarray(kf["kps"]), np.ascontiguousarray(kf["desc"], np.uint8),
                 np.ascontiguousarray(kf["has_mp"], np.uint8), np.ascontiguousarray(kf["node_ids"], np.uint32),
                 np.ascontiguousarray(kf["node_start"], np.int32), np.ascontiguousarray(kf["node_feats"], np.uint32)]
            keep.append(a)
            return KeyFrameFeatures(len(a[0]), len(a[3]), 0, _ptr(a[0]), _ptr(a[1]), None, _ptr(a[2]), _ptr(a[3]), _ptr(a[4]),
                                    _ptr(a[5]), None, None)
        v1, v2 = view(kf1), view(second)
        n_out = v2.n if second_is_frame else v1.n
        m = np.zeros(max(n_out, 1), np.int32); nm = np.zeros(1, np.int32)
        check(lib().ygzorb_search_by_bow(self.h, C.byref(v1), C.byref(v2), int(second_is_frame), C.c_float(nnratio),
                                         int(check_orientation), 0, _ptr(m), _ptr(nm)))
        return int(nm[0]), m[:n_out].copy()

    def stereo_match(self, ex_left, ex_right, frame, kps_l, desc_l, kps_r, desc_r, bf, b):
        kps_l, kps_r = np.ascontiguousarray(kps_l), np.ascontiguousarray(kps_r)
        desc_l, desc_r = np.ascontiguousarray(desc_l, np.uint8), np.ascontiguousarray(desc_r, np.uint8)
        n = len(kps_l)
        ur = np.zeros(max(n, 1), np.float32); dp = np.zeros(max(n, 1), np.float32)
        check(lib().ygzorb_stereo_match(self.h, ex_left.h, ex_right.h, frame, _ptr(kps_l), _ptr(desc_l), n, _ptr(kps_r),
                                        _ptr(desc_r), len(kps_r), bf, b, 0, _ptr(ur), _ptr(dp)))
        return ur[:n].copy(), dp[:n].copy()


class Vocabulary:
    """Device-resident DBoW2 vocabulary tree (include/ygzorb.h, bag-of-words section)."""

    def __init__(self, handle):
        self.h = handle
        info = np.zeros(6, np.int32)
        check(lib().ygzorb_vocabulary_info(self.h, _ptr(info)))
        self.n_nodes, self.k, self.L, self.n_words, self.n_children = [int(x) for x in info[:5]]

    @classmethod
    def from_arrays(cls, k, L, child_start, child_list, desc, word_id, weight, device=0):
        a = [np.ascontiguousarray(child_start, np.int32), np.ascontiguousarray(child_list, np.int32),
             np.ascontiguousarray(desc, np.uint8), np.ascontiguousarray(word_id, np.int32), np.ascontiguousarray(weight, np.float64)]
        h = C.c_void_p()
        check(lib().ygzorb_vocabulary_create(len(a[3]), int(k), int(L), *[C.c_void_p(_ptr(x)) for x in a], device, C.byref(h)))
        return cls(h)

    @classmethod
    def load_binary(cls, path, device=0):
        h = C.c_void_p()
        check(lib().ygzorb_vocabulary_load_binary(path.encode(), device, C.byref(h)))
        return cls(h)

    @classmethod
    def load_text(cls, path, device=0):
        h = C.c_void_p()
        check(lib().ygzorb_vocabulary_load_text(path.encode(), device, C.byref(h)))
        return cls(h)

    def arrays(self):
        n = self.n_nodes
        out = {"child_start": np.zeros(n + 1, np.int32), "child_list": np.zeros(max(self.n_children, 1), np.int32),
               "desc": np.zeros((n, 32), np.uint8), "word_id": np.zeros(n, np.int32), "weight": np.zeros(n, np.float64)}
        check(lib().ygzorb_vocabulary_arrays(self.h, *[C.c_void_p(_ptr(out[k])) for k in
                                                       ("child_start", "child_list", "desc", "word_id", "weight")]))
        out["child_list"] = out["child_list"][:self.n_children]
        return out

    def close(self):
        if self.h:
            lib().ygzorb_vocabulary_destroy(self.h)
            self.h = None


def bow_transform(matcher, voc, frames, levelsup, stride=None):
    """frames: list of (n_i, 32) uint8 descriptor arrays.  Returns per frame a dict with the keys word, node, weight,
    bow_word, bow_value, fv_node, fv_start, fv_feat."""
    nf = len(frames)
    st = stride or max(max((len(f) for f in frames), default=1), 1)
    desc = np.zeros((nf, st, 32), np.uint8)
    n_feat = np.array([len(f) for f in frames], np.int32)
    for i, f in enumerate(frames):
        desc[i, :len(f)] = f
    word = np.zeros((nf, st), np.int32); node = np.zeros((nf, st), np.int32); weight = np.zeros((nf, st), np.float64)
    bc = np.zeros(nf, np.int32); bw = np.zeros((nf, st), np.int32); bv = np.zeros((nf, st), np.float64)
    fc = np.zeros(nf, np.int32); fn = np.zeros((nf, st), np.uint32); fs = np.zeros((nf, st + 1), np.int32); ff = np.zeros((nf, st), np.uint32)
    check(lib().ygzorb_bow_transform(matcher.h, voc.h, C.c_void_p(_ptr(desc)), C.c_void_p(_ptr(n_feat)), nf, st, int(levelsup), 0,
                                     *[C.c_void_p(_ptr(a)) for a in (word, node, weight, bc, bw, bv, fc, fn, fs, ff)]))
    out = []
    for i in range(nf):
        n, nb, nn = int(n_feat[i]), int(bc[i]), int(fc[i])
        out.append({"word": word[i, :n].copy(), "node": node[i, :n].copy(), "weight": weight[i, :n].copy(),
                    "bow_word": bw[i, :nb].copy(), "bow_value": bv[i, :nb].copy(), "fv_node": fn[i, :nn].astype(np.int32),
                    "fv_start": fs[i, :nn + 1].copy(), "fv_feat": ff[i, :int(fs[i, nn])].astype(np.int32)})
    return out


BROADCAST_TRAIN = 16   # YGZORB_BROADCAST_TRAIN


def shard_range(n, rank, world):
    """ygzorb_shard_range: contiguous [begin, end) of rank `rank` (host logic, no device needed)."""
    b, e = C.c_longlong(), C.c_longlong()
    check(lib().ygzorb_shard_range(C.c_longlong(n), int(rank), int(world), C.byref(b), C.byref(e)))
    return b.value, e.value


class Comm:
    """ygzorb_comm: the gather layer of the two sharded workloads (frames of a batched extraction, queries of a brute-force kNN)."""

    def __init__(self, rank=0, world=1, device=0, unique_id=None):
        L = lib()
        L.ygzorb_comm_destroy.restype = None
        L.ygzorb_comm_destroy.argtypes = [C.c_void_p]
        L.ygzorb_comm_launch_count.restype = C.c_longlong
        L.ygzorb_comm_launch_count.argtypes = [C.c_void_p]
        L.ygzorb_comm_create_rank.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.ygzorb_hamming_knn2_sharded.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p, C.c_int, C.c_float,
                                                  C.c_int, C.c_uint, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ygzorb_extract_batch_sharded.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_int, C.c_int,
                                                   C.c_size_t, C.c_uint, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        self.rank, self.world, self.device = rank, world, device
        h = C.c_void_p()
        idp = None
        if world > 1:
            self._id = (C.c_uint8 * 128).from_buffer_copy(bytes(unique_id))
            idp = C.cast(self._id, C.c_void_p)
        check(L.ygzorb_comm_create_rank(idp, rank, world, device, C.byref(h)))
        self.h = h
        _track(self)

    @staticmethod
    def unique_id():
        buf = (C.c_uint8 * 128)()
        check(lib().ygzorb_comm_unique_id(buf))
        return bytes(buf)

    @classmethod
    def from_torch_distributed(cls, device):
        """One process per GPU under torchrun: rank 0 makes the NCCL id, torch.distributed carries the 128 bytes."""
        import torch
        import torch.distributed as dist
        rank, world = dist.get_rank(), dist.get_world_size()
        t = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            t = torch.frombuffer(bytearray(cls.unique_id()), dtype=torch.uint8).clone()
        t = t.to("cuda:%d" % device)
        dist.broadcast(t, src=0)
        return cls(rank, world, device, bytes(t.cpu().numpy().tobytes()))

    def close(self):
        if getattr(self, "h", None):
            lib().ygzorb_comm_destroy(self.h)
            self.h = None

    __del__ = close

    def launch_count(self):
        return int(lib().ygzorb_comm_launch_count(self.h))

    def hamming_knn2_sharded(self, matcher, q_ptr, nq_total, t_ptr, nt, ratio, th, flags, idx_ptr, d1_ptr, d2_ptr, match_ptr):
        check(lib().ygzorb_hamming_knn2_sharded(self.h, matcher.h, q_ptr, nq_total, t_ptr, nt, ratio, th, flags, idx_ptr, d1_ptr,
                                                d2_ptr, match_ptr))

    def extract_batch_sharded(self, ex, frames_ptr, n_total, w, h, stride, frame_stride, flags, cap, chunk, counts_ptr, off_ptr,
                              kps_ptr, desc_ptr):
        check(lib().ygzorb_extract_batch_sharded(self.h, ex.h, frames_ptr, n_total, w, h, stride, frame_stride, flags, cap, chunk,
                                                 counts_ptr, off_ptr, kps_ptr, desc_ptr))


NOALLOC_SCRATCH_BYTES = 8 * 1024 * 1024   # YGZORB_NOALLOC_SCRATCH_BYTES
NOALLOC_MAX_QUERIES = 600000              # YGZORB_NOALLOC_MAX_QUERIES


def hamming_knn2_noalloc(device, stream, q_ptr, nq, t_ptr, nt, ratio, th, idx_ptr, d1_ptr, d2_ptr, match_ptr):
    """ygzorb_hamming_knn2_noalloc: device pointers only, enqueued on `stream` (int cudaStream_t or None), not synchronised."""
    check(lib().ygzorb_hamming_knn2_noalloc(device, stream, q_ptr, nq, t_ptr, nt, ratio, th, idx_ptr, d1_ptr, d2_ptr, match_ptr))


def hamming_knn2(query, train, ratio, th, device=0):
    m = Matcher(device)
    try:
        return m.hamming_knn2(query, train, ratio, th)
    finally:
        m.close()
