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
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


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


def remap(img, map1, map2):
    img, w, h = _img(img)
    m1 = np.ascontiguousarray(map1, np.int16); m2 = np.ascontiguousarray(map2, np.uint16)
    dh, dw = m2.shape
    out = np.zeros((dh, dw), np.uint8)
    lib().ygzo_remap_u8(_p(img), w, h, w, _p(m1), _p(m2), dw, dh, _p(out))
    return out


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


# ---- part 2: matchers (oracle/match_oracle.cpp) ------------------------------------------------
def descriptor_distance(a, b):
    a = np.ascontiguousarray(a, np.uint8)
    b = np.ascontiguousarray(b, np.uint8)
    return lib().ygzo_descriptor_distance(_p(a), _p(b))


def distinctive_descriptors(obs_desc, obs_start, best_desc=None):
    obs_desc = np.ascontiguousarray(obs_desc, np.uint8).reshape(-1, 32)
    obs_start = np.ascontiguousarray(obs_start, np.int32)
    n = len(obs_start) - 1
    idx, med = np.zeros(n, np.int32), np.zeros(n, np.int32)
    out = np.zeros((n, 32), np.uint8) if best_desc is None else np.ascontiguousarray(best_desc, np.uint8)
    lib().ygzo_distinctive_descriptors(_p(obs_desc), _p(obs_start), n, _p(idx), _p(med), _p(out))
    return idx, med, out


class DirectKeyFrame(C.Structure):
    _fields_ = [("levels", C.c_void_p * 8), ("w", C.c_int32 * 8), ("h", C.c_int32 * 8), ("pitch", C.c_int32 * 8),
                ("pose_ref", C.c_float * 7), ("t_cur_ref", C.c_float * 7),
                ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float)]


def find_direct_projection(kfs, cur_levels, cur_intr, scale, inv_scale, inv_sigma2_1, kf_index, px_ref, octave, world_pos, px_curr):
    """kfs: list of dicts {levels: [images], pose_ref, t_cur_ref, fx, fy, cx, cy}; cur_levels: list of images."""
    keep = []
    arr = (DirectKeyFrame * max(len(kfs), 1))()
    for a, k in zip(arr, kfs):
        imgs = [np.ascontiguousarray(x, np.uint8) for x in k["levels"]]
        keep.append(imgs)
        for l, im in enumerate(imgs):
            a.levels[l] = im.ctypes.data; a.w[l] = im.shape[1]; a.h[l] = im.shape[0]; a.pitch[l] = im.shape[1]
        a.pose_ref[:] = [float(x) for x in k["pose_ref"]]
        a.t_cur_ref[:] = [float(x) for x in k["t_cur_ref"]]
        a.fx, a.fy, a.cx, a.cy = (float(k[n]) for n in ("fx", "fy", "cx", "cy"))
    cur = [np.ascontiguousarray(x, np.uint8) for x in cur_levels]
    ptrs = (C.c_void_p * len(cur))(*[x.ctypes.data for x in cur])
    cw = np.array([x.shape[1] for x in cur], np.int32); ch = np.array([x.shape[0] for x in cur], np.int32)
    kf_index = np.ascontiguousarray(kf_index, np.int32); octave = np.ascontiguousarray(octave, np.int32)
    px_ref = np.ascontiguousarray(px_ref, np.float32); world_pos = np.ascontiguousarray(world_pos, np.float32)
    out = np.array(px_curr, np.float32).reshape(-1, 2).copy()
    n = len(kf_index)
    intr = np.ascontiguousarray(cur_intr, np.float32)
    scale = np.ascontiguousarray(scale, np.float32); inv_scale = np.ascontiguousarray(inv_scale, np.float32)
    level, ok = np.zeros(n, np.int32), np.zeros(n, np.uint8)
    lib().ygzo_find_direct_projection(arr, len(kfs), ptrs, _p(cw), _p(ch), _p(cw), len(cur), _p(intr), _p(scale), _p(inv_scale),
                                      C.c_float(inv_sigma2_1), n, _p(kf_index), _p(px_ref), _p(octave), _p(world_pos), _p(out),
                                      _p(level), _p(ok))
    return out, level, ok


class OPyramid(C.Structure):
    _fields_ = [("levels", C.c_void_p * 8), ("w", C.c_int32 * 8), ("h", C.c_int32 * 8), ("pitch", C.c_int32 * 8)]


def _opyr(levels):
    imgs = [np.ascontiguousarray(x, np.uint8) for x in levels]
    P = OPyramid()
    for l, im in enumerate(imgs):
        P.levels[l] = im.ctypes.data; P.w[l] = im.shape[1]; P.h[l] = im.shape[0]; P.pitch[l] = im.shape[1]
    return P, imgs


def sparse_align_residuals(ref_levels, cur_levels, level, scale, focal, cur_intr, ref_pose, t_cur_ref, kp_xy, valid, world_pos,
                           accumulate=0):
    R, k1 = _opyr(ref_levels); Cu, k2 = _opyr(cur_levels)
    kp_xy = np.ascontiguousarray(kp_xy, np.float32); valid = np.ascontiguousarray(valid, np.uint8)
    world_pos = np.ascontiguousarray(world_pos, np.float32)
    n = len(valid)
    H, J, chi2, nm = np.zeros((6, 6), np.float32), np.zeros(6, np.float32), C.c_float(), C.c_int64()
    vis, patch, jac = np.zeros(n, np.uint8), np.zeros((n, 16), np.float32), np.zeros((n, 16, 6), np.float32)
    lib().ygzo_sparse_align_residuals(C.byref(R), C.byref(Cu), level, C.c_float(scale), C.c_float(focal),
                                      _p(np.ascontiguousarray(cur_intr, np.float32)), _p(np.ascontiguousarray(ref_pose, np.float32)),
                                      _p(np.ascontiguousarray(t_cur_ref, np.float32)), n, _p(kp_xy), _p(valid), _p(world_pos),
                                      accumulate, _p(H), _p(J), C.byref(chi2), C.byref(nm), _p(vis), _p(patch), _p(jac))
    return H, J, chi2.value, nm.value, vis, patch, jac


def sparse_align_run(ref_levels, cur_levels, max_level, min_level, inv_scale, focal, cur_intr, ref_pose, kp_xy, valid, world_pos,
                     t_cur_ref, accumulate=0):
    R, k1 = _opyr(ref_levels); Cu, k2 = _opyr(cur_levels)
    kp_xy = np.ascontiguousarray(kp_xy, np.float32); valid = np.ascontiguousarray(valid, np.uint8)
    world_pos = np.ascontiguousarray(world_pos, np.float32)
    T = np.array(t_cur_ref, np.float32).copy()
    stats = np.zeros(2, np.float32)
    lib().ygzo_sparse_align_run.restype = C.c_int64
    n = lib().ygzo_sparse_align_run(C.byref(R), C.byref(Cu), max_level, min_level, _p(np.ascontiguousarray(inv_scale, np.float32)),
                                    C.c_float(focal), _p(np.ascontiguousarray(cur_intr, np.float32)),
                                    _p(np.ascontiguousarray(ref_pose, np.float32)), len(valid), _p(kp_xy), _p(valid), _p(world_pos),
                                    accumulate, _p(T), _p(stats))
    return T, n, stats


def hamming_knn2(query, train, ratio, th, threads=1):
    query = np.ascontiguousarray(query, np.uint8)
    train = np.ascontiguousarray(train, np.uint8)
    nq, nt = len(query), len(train)
    idx, d1, d2, match = (np.zeros(nq, np.int32) for _ in range(4))
    lib().ygzo_hamming_knn2(_p(query), nq, _p(train), nt, C.c_float(ratio), th, threads, _p(idx), _p(d1), _p(d2), _p(match))
    return idx, d1, d2, match


def grid_inv(min_x, min_y, max_x, max_y):
    """mfGridElementWidthInv / HeightInv (src/Frame.cc:296-297), float arithmetic."""
    inv_w = np.float32(64) / (np.float32(max_x) - np.float32(min_x))
    inv_h = np.float32(48) / (np.float32(max_y) - np.float32(min_y))
    return np.float32(inv_w), np.float32(inv_h)


def search_by_projection(kps, desc, u_right, claimed, bounds, scale_factors, mp, th, check_level, nnratio):
    """mp: dict with proj_x, proj_y, proj_xr, view_cos (f4), level (i4), in_view, has_obs (u1), desc (m,32)."""
    n, m = len(kps), len(mp["proj_x"])
    kps = np.ascontiguousarray(kps)
    desc = np.ascontiguousarray(desc, np.uint8)
    min_x, min_y, max_x, max_y = bounds
    inv_w, inv_h = grid_inv(*bounds)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    kp_to_mp = np.zeros(max(n, 1), np.int32)
    mp_best = np.zeros(max(m, 1), np.int32)
    f32 = lambda a: np.ascontiguousarray(a, np.float32)
    u8 = lambda a: np.ascontiguousarray(a, np.uint8)
    px, py, pxr, vc = f32(mp["proj_x"]), f32(mp["proj_y"]), f32(mp["proj_xr"]), f32(mp["view_cos"])
    lv, iv, ho, md = np.ascontiguousarray(mp["level"], np.int32), u8(mp["in_view"]), u8(mp["has_obs"]), u8(mp["desc"])
    ur = f32(u_right) if u_right is not None else None
    cl = u8(claimed) if claimed is not None else None
    nm = lib().ygzo_search_by_projection(_p(kps), _p(desc), _p(ur), _p(cl), n, C.c_float(min_x), C.c_float(min_y),
                                         C.c_float(inv_w), C.c_float(inv_h), _p(sf), _p(px), _p(py), _p(pxr), _p(vc),
                                         _p(lv), _p(iv), _p(ho), _p(md), m, C.c_float(th), int(check_level),
                                         C.c_float(nnratio), _p(kp_to_mp), _p(mp_best))
    return nm, kp_to_mp[:n].copy(), mp_best[:m].copy()


def search_by_projection_last(kps, desc, u_right, claimed, bounds, scale_factors, last, th, mono, check_level,
                              check_orientation):
    """last: dict with valid, has_obs (u1), world_pos (m,3) f4, desc (m,32), octave (i4), angle (f4), pose (30,) f4."""
    n, m = len(kps), len(last["valid"])
    kps = np.ascontiguousarray(kps)
    desc = np.ascontiguousarray(desc, np.uint8)
    inv_w, inv_h = grid_inv(*bounds)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    f32 = lambda a: np.ascontiguousarray(a, np.float32)
    u8 = lambda a: np.ascontiguousarray(a, np.uint8)
    ur = f32(u_right) if u_right is not None else None
    cl = u8(claimed) if claimed is not None else None
    va, ho, wp, pd = u8(last["valid"]), u8(last["has_obs"]), f32(last["world_pos"]), u8(last["desc"])
    oc, an, pose = np.ascontiguousarray(last["octave"], np.int32), f32(last["angle"]), f32(last["pose"])
    kp_to_pt = np.zeros(max(n, 1), np.int32)
    best = np.zeros(max(m, 1), np.int32)
    nm = lib().ygzo_search_by_projection_last(_p(kps), _p(desc), _p(ur), _p(cl), n, C.c_float(bounds[0]), C.c_float(bounds[1]),
                                              C.c_float(bounds[2]), C.c_float(bounds[3]), C.c_float(inv_w), C.c_float(inv_h),
                                              _p(sf), _p(pose), _p(va), _p(ho), _p(wp), _p(pd), _p(oc), _p(an), m,
                                              C.c_float(th), int(mono), int(check_level), int(check_orientation),
                                              _p(kp_to_pt), _p(best))
    return nm, kp_to_pt[:n].copy(), best[:m].copy()


def search_by_projection_keyframe(kps, desc, claimed, bounds, scale_factors, pts, th, variant, orb_dist=100,
                                  check_orientation=False, u_right=None, inv_sigma2=None):
    """variant: 0 relocalisation, 1 Sim3 SearchByProjection, 2 Fuse, 3 Fuse with Sim3, 4 one direction of SearchBySim3
    (False / True select 0 / 1).  pts: dict as tests/synth.keyframe_points_case makes it.
    Returns (n_matches, kp_to_pt, pt_best_idx)."""
    n, m = len(kps), len(pts["valid"])
    kps = np.ascontiguousarray(kps)
    desc = np.ascontiguousarray(desc, np.uint8)
    inv_w, inv_h = grid_inv(*bounds)
    f32 = lambda a: np.ascontiguousarray(a, np.float32)
    u8 = lambda a: np.ascontiguousarray(a, np.uint8)
    sf = f32(scale_factors)
    cl = u8(claimed) if claimed is not None else None
    ur = f32(u_right) if u_right is not None else None
    is2 = f32(inv_sigma2) if inv_sigma2 is not None else None
    va, wp, nr, pd, an = u8(pts["valid"]), f32(pts["world_pos"]), f32(pts["normal"]), u8(pts["desc"]), f32(pts["angle"])
    mx, mn, pose = f32(pts["max_distance"]), f32(pts["min_distance"]), f32(pts["pose"])
    kp_to_pt = np.zeros(max(n, 1), np.int32)
    best = np.zeros(max(m, 1), np.int32)
    nm = lib().ygzo_search_by_projection_keyframe(
        _p(kps), _p(desc), _p(cl), _p(ur), n, C.c_float(bounds[0]), C.c_float(bounds[1]), C.c_float(bounds[2]),
        C.c_float(bounds[3]), C.c_float(inv_w), C.c_float(inv_h), _p(sf), _p(is2), _p(pose), _p(va), _p(wp), _p(nr), _p(pd),
        _p(an), _p(mx), _p(mn), m, C.c_float(pts["log_scale_factor"]), int(pts["n_scale_levels"]), C.c_float(th),
        int(orb_dist), int(check_orientation), int(variant), _p(kp_to_pt), _p(best))
    return nm, kp_to_pt[:n].copy(), best[:m].copy()


def search_for_initialization(kps1, desc1, kps2, desc2, bounds2, prev_matched, window_size, nnratio, check_orientation):
    """Returns (n_matches, vnMatches12, updated vbPrevMatched)."""
    k1, k2 = np.ascontiguousarray(kps1), np.ascontiguousarray(kps2)
    d1, d2 = np.ascontiguousarray(desc1, np.uint8), np.ascontiguousarray(desc2, np.uint8)
    inv_w, inv_h = grid_inv(*bounds2)
    pm = np.array(prev_matched, np.float32).reshape(-1, 2).copy()
    m12 = np.zeros(max(len(k1), 1), np.int32)
    nm = lib().ygzo_search_for_initialization(_p(k1), _p(d1), len(k1), _p(k2), _p(d2), len(k2), C.c_float(bounds2[0]),
                                              C.c_float(bounds2[1]), C.c_float(inv_w), C.c_float(inv_h), _p(pm), int(window_size),
                                              C.c_float(nnratio), int(check_orientation), _p(m12))
    return nm, m12[:len(k1)].copy(), pm


def sim3_agreement(match1, match2):
    m1, m2 = np.ascontiguousarray(match1, np.int32), np.ascontiguousarray(match2, np.int32)
    out = np.zeros(max(len(m1), 1), np.int32)
    n = lib().ygzo_sim3_agreement(_p(m1), len(m1), _p(m2), len(m2), _p(out))
    return n, out[:len(m1)].copy()


def search_by_bow(kf1, second, second_is_frame, nnratio, check_orientation):
    """kf dicts: kps, desc, has_mp, node_ids, node_start, node_feats.  Returns (n_matches, matches)."""
    def pk(kf):
        return (np.ascontiguousarray(kf["kps"]), np.ascontiguousarray(kf["desc"], np.uint8), np.ascontiguousarray(kf["has_mp"], np.uint8),
                np.ascontiguousarray(kf["node_ids"], np.uint32), np.ascontiguousarray(kf["node_start"], np.int32),
                np.ascontiguousarray(kf["node_feats"], np.uint32))
    a, b = pk(kf1), pk(second)
    n_out = len(b[0]) if second_is_frame else len(a[0])
    m = np.zeros(max(n_out, 1), np.int32)
    nm = lib().ygzo_search_by_bow(_p(a[0]), _p(a[1]), _p(a[2]), len(a[0]), _p(a[3]), _p(a[4]), _p(a[5]), len(a[3]),
                                  _p(b[0]), _p(b[1]), _p(b[2]), len(b[0]), _p(b[3]), _p(b[4]), _p(b[5]), len(b[3]),
                                  int(second_is_frame), C.c_float(nnratio), int(check_orientation), _p(m))
    return nm, m[:n_out].copy()


def search_for_triangulation(kf1, kf2, F12, ex, ey, only_stereo, check_orientation):
    """kf: dict with kps, desc, u_right, has_mp, node_ids, node_start, node_feats, scale_factors, level_sigma2."""
    def pk(kf):
        return (np.ascontiguousarray(kf["kps"]), np.ascontiguousarray(kf["desc"], np.uint8),
                np.ascontiguousarray(kf["u_right"], np.float32), np.ascontiguousarray(kf["has_mp"], np.uint8),
                np.ascontiguousarray(kf["node_ids"], np.uint32), np.ascontiguousarray(kf["node_start"], np.int32),
                np.ascontiguousarray(kf["node_feats"], np.uint32))
    a, b = pk(kf1), pk(kf2)
    sf2 = np.ascontiguousarray(kf2["scale_factors"], np.float32)
    s2 = np.ascontiguousarray(kf2["level_sigma2"], np.float32)
    F = np.ascontiguousarray(F12, np.float32).reshape(9)
    m12 = np.zeros(max(len(a[0]), 1), np.int32)
    nm = lib().ygzo_search_for_triangulation(
        _p(a[0]), _p(a[1]), _p(a[2]), _p(a[3]), len(a[0]), _p(a[4]), _p(a[5]), _p(a[6]), len(a[4]),
        _p(b[0]), _p(b[1]), _p(b[2]), _p(b[3]), len(b[0]), _p(b[4]), _p(b[5]), _p(b[6]), len(b[4]),
        _p(sf2), _p(s2), _p(F), C.c_float(ex), C.c_float(ey), int(only_stereo), int(check_orientation), _p(m12))
    return nm, m12[:len(a[0])].copy()


def stereo_match(kps_l, desc_l, kps_r, desc_r, pyr_l, pyr_r, scale_factors, inv_scale_factors, bf, b):
    """pyr_*: list of level images (uint8 2-D).  Returns (u_right, depth, best_sad)."""
    kps_l, kps_r = np.ascontiguousarray(kps_l), np.ascontiguousarray(kps_r)
    desc_l, desc_r = np.ascontiguousarray(desc_l, np.uint8), np.ascontiguousarray(desc_r, np.uint8)
    lw = np.array([l.shape[1] for l in pyr_l], np.int32)
    lh = np.array([l.shape[0] for l in pyr_l], np.int32)
    off = np.zeros(len(pyr_l), np.int64)
    off[1:] = np.cumsum(lw[:-1].astype(np.int64) * lh[:-1])
    pl = np.concatenate([np.ascontiguousarray(l, np.uint8).ravel() for l in pyr_l])
    pr = np.concatenate([np.ascontiguousarray(l, np.uint8).ravel() for l in pyr_r])
    n = len(kps_l)
    ur, dp = np.zeros(max(n, 1), np.float32), np.zeros(max(n, 1), np.float32)
    bd = np.zeros(max(n, 1), np.int32)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    isf = np.ascontiguousarray(inv_scale_factors, np.float32)
    lib().ygzo_stereo_match(_p(kps_l), _p(desc_l), n, _p(kps_r), _p(desc_r), len(kps_r), _p(pl), _p(pr), _p(off), _p(lw),
                            _p(lh), _p(sf), _p(isf), C.c_float(bf), C.c_float(b), _p(ur), _p(dp), _p(bd))
    return ur[:n].copy(), dp[:n].copy(), bd[:n].copy()


def search_by_projection_batch(n_kp, n_mp, kps, desc, bounds, scale_factors, mp, th, check_level, nnratio, threads=1):
    """Arrays at fixed strides: kps (F, ks) KP_DTYPE, desc (F, ks, 32), mp[...] (F, ms[,32]).  Returns (kp_to_mp, n_matches)."""
    F, ks = kps.shape
    ms = mp["proj_x"].shape[1]
    inv_w, inv_h = grid_inv(*bounds)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    kp_to_mp = np.zeros((F, ks), np.int32)
    nm = np.zeros(F, np.int32)
    lib().ygzo_search_by_projection_batch(
        F, ks, ms, _p(n_kp), _p(n_mp), _p(kps), _p(desc), C.c_float(bounds[0]), C.c_float(bounds[1]), C.c_float(inv_w),
        C.c_float(inv_h), _p(sf), _p(mp["proj_x"]), _p(mp["proj_y"]), _p(mp["proj_xr"]), _p(mp["view_cos"]), _p(mp["level"]),
        _p(mp["in_view"]), _p(mp["has_obs"]), _p(mp["desc"]), C.c_float(th), int(check_level), C.c_float(nnratio), threads,
        _p(kp_to_mp), _p(nm))
    return kp_to_mp, nm


# ---- bag of words (oracle/bow_oracle.cpp)
class Vocabulary:
    """Restated DBoW2 vocabulary tree (oracle side)."""

    def __init__(self, handle):
        if not handle:
            raise RuntimeError("vocabulary could not be loaded")
        self.h = C.c_void_p(handle)
        info = np.zeros(5, np.int32)
        lib().ygzo_voc_info(self.h, _p(info))
        self.n_nodes, self.k, self.L, self.n_words, self.n_children = [int(x) for x in info]

    @classmethod
    def load_binary(cls, path):
        lib().ygzo_voc_load_binary.restype = C.c_void_p
        return cls(lib().ygzo_voc_load_binary(path.encode()))

    @classmethod
    def from_arrays(cls, k, L, child_start, child_list, desc, word_id, weight):
        lib().ygzo_voc_from_arrays.restype = C.c_void_p
        a = [np.ascontiguousarray(child_start, np.int32), np.ascontiguousarray(child_list, np.int32),
             np.ascontiguousarray(desc, np.uint8), np.ascontiguousarray(word_id, np.int32), np.ascontiguousarray(weight, np.float64)]
        return cls(lib().ygzo_voc_from_arrays(len(a[3]), k, L, *[_p(x) for x in a]))

    def arrays(self):
        n = self.n_nodes
        out = {"parent": np.zeros(n, np.int32), "child_start": np.zeros(n + 1, np.int32),
               "child_list": np.zeros(max(self.n_children, 1), np.int32), "desc": np.zeros((n, 32), np.uint8),
               "word_id": np.zeros(n, np.int32), "weight": np.zeros(n, np.float64)}
        lib().ygzo_voc_arrays(self.h, _p(out["parent"]), _p(out["child_start"]), _p(out["child_list"]), _p(out["desc"]),
                              _p(out["word_id"]), _p(out["weight"]))
        out["child_list"] = out["child_list"][:self.n_children]
        return out

    def transform(self, desc, levelsup):
        """Returns dict: word, node (n,), weight (n,) f8; bow_word, bow_value; fv_node, fv_start, fv_feat (CSR)."""
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = len(desc)
        word = np.zeros(max(n, 1), np.int32); node = np.zeros(max(n, 1), np.int32); weight = np.zeros(max(n, 1), np.float64)
        bw = np.zeros(max(n, 1), np.int32); bv = np.zeros(max(n, 1), np.float64)
        nn = np.zeros(1, np.int32)
        fn = np.zeros(max(n, 1), np.int32); fs = np.zeros(n + 1, np.int32); ff = np.zeros(max(n, 1), np.int32)
        nb = lib().ygzo_bow_transform(self.h, _p(desc), n, int(levelsup), _p(word), _p(node), _p(weight), _p(bw), _p(bv), _p(nn),
                                      _p(fn), _p(fs), _p(ff))
        k = int(nn[0])
        return {"word": word[:n], "node": node[:n], "weight": weight[:n], "bow_word": bw[:nb].copy(), "bow_value": bv[:nb].copy(),
                "fv_node": fn[:k].copy(), "fv_start": fs[:k + 1].copy(), "fv_feat": ff[:int(fs[k])].copy()}

    def close(self):
        if self.h:
            lib().ygzo_voc_free(self.h)
            self.h = None


REF_BOW = os.path.join(ROOT, "oracle", "_ref", "ygz_ref_bow")
