"""ctypes binding of oracle/_ref/libygz_ref_match.so: the REFERENCE'S OWN src/ORBmatcher.cc, src/Align.cc and the hot-path
member functions of Frame / KeyFrame / MapPoint, compiled unmodified (oracle/Makefile, target refmatch; oracle/ref_match_api.cpp).
Test infrastructure.  Same Python signatures as the restatement's binding in oracle_lib.py, so a test reads
`assert O.search_x(...) == R.search_x(...)`.  The library only exists where the reference tree was present at build time (the
authoring container; it travels to the GPU box with the snapshot): tests skip on `not available()` and fall back to the
committed golden vectors this library produced (tests/golden/match_*.npz, tools/make_golden_match.py)."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(ROOT, "oracle", "_ref", "libygz_ref_match.so")
_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(LIB_PATH)
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


f32 = lambda a: np.ascontiguousarray(a, np.float32)
u8 = lambda a: np.ascontiguousarray(a, np.uint8)
i32 = lambda a: np.ascontiguousarray(a, np.int32)
u32 = lambda a: np.ascontiguousarray(a, np.uint32)
cf = C.c_float


def grid_inv(min_x, min_y, max_x, max_y):
    inv_w = np.float32(64) / (np.float32(max_x) - np.float32(min_x))
    inv_h = np.float32(48) / (np.float32(max_y) - np.float32(min_y))
    return np.float32(inv_w), np.float32(inv_h)


def descriptor_distance(a, b):
    return lib().ygzr_descriptor_distance(_p(u8(a)), _p(u8(b)))


def search_by_projection(kps, desc, u_right, claimed, bounds, scale_factors, mp, th, check_level, nnratio):
    """ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th, checkLevel).  Returns (n_matches, kp_to_mp)."""
    n, m = len(kps), len(mp["proj_x"])
    kps, desc = np.ascontiguousarray(kps), u8(desc)
    inv_w, inv_h = grid_inv(*bounds)
    ur = f32(u_right) if u_right is not None else None
    cl = u8(claimed) if claimed is not None else None
    out = np.zeros(max(n, 1), np.int32)
    a = [f32(mp["proj_x"]), f32(mp["proj_y"]), f32(mp["proj_xr"]), f32(mp["view_cos"]), i32(mp["level"]), u8(mp["in_view"]),
         u8(mp["has_obs"]), u8(mp["desc"])]
    sf = f32(scale_factors)
    nm = lib().ygzr_search_by_projection(_p(kps), _p(desc), _p(ur), _p(cl), n, cf(bounds[0]), cf(bounds[1]), cf(inv_w), cf(inv_h),
                                         _p(sf), *[_p(x) for x in a], m, cf(th), int(check_level), cf(nnratio), _p(out))
    return nm, out[:n].copy()


def search_by_projection_last(kps, desc, u_right, claimed, bounds, scale_factors, last, th, mono, check_level, check_orientation):
    n, m = len(kps), len(last["valid"])
    kps, desc = np.ascontiguousarray(kps), u8(desc)
    inv_w, inv_h = grid_inv(*bounds)
    ur = f32(u_right) if u_right is not None else None
    cl = u8(claimed) if claimed is not None else None
    out = np.zeros(max(n, 1), np.int32)
    a = [f32(scale_factors), f32(last["pose"]), u8(last["valid"]), u8(last["has_obs"]), f32(last["world_pos"]), u8(last["desc"]),
         i32(last["octave"]), f32(last["angle"])]
    nm = lib().ygzr_search_by_projection_last(_p(kps), _p(desc), _p(ur), _p(cl), n, cf(bounds[0]), cf(bounds[1]), cf(bounds[2]),
                                              cf(bounds[3]), cf(inv_w), cf(inv_h), *[_p(x) for x in a], m, cf(th), int(mono),
                                              int(check_level), int(check_orientation), _p(out))
    return nm, out[:n].copy()


def decompose_scw(scw):
    """Scw (4x4) -> (Rcw 9, tcw 3, Ow 3) by the cv::Mat expressions of src/ORBmatcher.cc:275-278."""
    scw = f32(scw).reshape(16)
    out = np.zeros(15, np.float32)
    lib().ygzr_decompose_scw(_p(scw), _p(out))
    return out


def scw_of_pose(pose, scale=1.0):
    """A Sim3 matrix [s R | s t; 0 0 0 1] from a case's Rcw (pose[0:9]) / tcw (pose[9:12])."""
    S = np.eye(4, dtype=np.float32)
    S[:3, :3] = np.float32(scale) * f32(pose[:9]).reshape(3, 3)
    S[:3, 3] = np.float32(scale) * f32(pose[9:12])
    return S


def search_by_projection_keyframe(kps, desc, claimed, bounds, scale_factors, pts, th, variant, orb_dist=100,
                                  check_orientation=False, u_right=None, inv_sigma2=None, scw=None):
    """variant 0 relocalisation, 1 Sim3 SearchByProjection (needs scw), 2 Fuse, 3 Fuse with Sim3 (needs scw).
    Returns (n_matches, kp_to_pt, pt_best_idx) — pt_best_idx only for the Fuse variants (else None)."""
    n, m = len(kps), len(pts["valid"])
    kps, desc = np.ascontiguousarray(kps), u8(desc)
    inv_w, inv_h = grid_inv(*bounds)
    cl = u8(claimed) if claimed is not None else None
    ur = f32(u_right) if u_right is not None else None
    is2 = f32(inv_sigma2) if inv_sigma2 is not None else None
    sc = f32(scw).reshape(16) if scw is not None else None
    assert (variant in (1, 3)) == (sc is not None)
    a = [f32(pts["pose"]), u8(pts["valid"]), f32(pts["world_pos"]), f32(pts["normal"]), u8(pts["desc"]), f32(pts["angle"]),
         f32(pts["max_distance"]), f32(pts["min_distance"])]
    sf = f32(scale_factors)
    kp_to_pt = np.zeros(max(n, 1), np.int32)
    best = np.full(max(m, 1), -1, np.int32)
    nm = lib().ygzr_search_by_projection_keyframe(
        _p(kps), _p(desc), _p(cl), _p(ur), n, cf(bounds[0]), cf(bounds[1]), cf(bounds[2]), cf(bounds[3]), cf(inv_w), cf(inv_h),
        _p(sf), _p(is2), *[_p(x) for x in a], m, cf(pts["log_scale_factor"]), int(pts["n_scale_levels"]), cf(th), int(orb_dist),
        int(check_orientation), int(variant), _p(sc), _p(kp_to_pt), _p(best))
    return nm, kp_to_pt[:n].copy(), (best[:m].copy() if variant >= 2 else None)


def search_by_sim3(kps1, desc1, kps2, desc2, bounds, scale_factors, intr, pose1, pose2, pts1, pts2, s12, R12, t12, th,
                   already12=None):
    """ORBmatcher::SearchBySim3, whole.  ptsK: dict with valid, world_pos, desc, max_distance, min_distance (one MapPoint per
    keypoint of keyframe K), log_scale_factor, n_scale_levels.  Returns (n_found, matches12, derived) with derived =
    (sR12 9, sR21 9, t21 3) as src/ORBmatcher.cc:1022-1025 computes them."""
    n1, n2 = len(kps1), len(kps2)
    inv_w, inv_h = grid_inv(*bounds)
    k1, k2, d1, d2 = np.ascontiguousarray(kps1), np.ascontiguousarray(kps2), u8(desc1), u8(desc2)
    sf = f32(scale_factors)
    a1 = [u8(pts1["valid"]), f32(pts1["world_pos"]), u8(pts1["desc"]), f32(pts1["max_distance"]), f32(pts1["min_distance"])]
    a2 = [u8(pts2["valid"]), f32(pts2["world_pos"]), u8(pts2["desc"]), f32(pts2["max_distance"]), f32(pts2["min_distance"])]
    al = i32(already12) if already12 is not None else None
    out = np.zeros(max(n1, 1), np.int32)
    derived = np.zeros(21, np.float32)
    R, t, it, p1, p2 = f32(R12).reshape(9), f32(t12).reshape(3), f32(intr), f32(pose1), f32(pose2)
    nf = lib().ygzr_search_by_sim3(_p(k1), _p(d1), n1, _p(k2), _p(d2), n2, cf(bounds[0]), cf(bounds[1]), cf(bounds[2]), cf(bounds[3]),
                                   cf(inv_w), cf(inv_h), _p(sf), int(pts1["n_scale_levels"]), cf(pts1["log_scale_factor"]), _p(it),
                                   _p(p1), _p(p2), *[_p(x) for x in a1], *[_p(x) for x in a2], cf(s12), _p(R), _p(t), cf(th),
                                   _p(al), _p(out), _p(derived))
    return nf, out[:n1].copy(), derived


def search_for_initialization(kps1, desc1, kps2, desc2, bounds2, prev_matched, window_size, nnratio, check_orientation):
    k1, k2, d1, d2 = np.ascontiguousarray(kps1), np.ascontiguousarray(kps2), u8(desc1), u8(desc2)
    inv_w, inv_h = grid_inv(*bounds2)
    pm = np.array(prev_matched, np.float32).reshape(-1, 2).copy()
    m12 = np.zeros(max(len(k1), 1), np.int32)
    nm = lib().ygzr_search_for_initialization(_p(k1), _p(d1), len(k1), _p(k2), _p(d2), len(k2), cf(bounds2[0]), cf(bounds2[1]),
                                              cf(inv_w), cf(inv_h), _p(pm), int(window_size), cf(nnratio), int(check_orientation),
                                              _p(m12))
    return nm, m12[:len(k1)].copy(), pm


def _pk(kf, with_ur=False):
    a = [np.ascontiguousarray(kf["kps"]), u8(kf["desc"])]
    if with_ur:
        a.append(f32(kf["u_right"]))
    return a + [u8(kf["has_mp"]), u32(kf["node_ids"]), i32(kf["node_start"]), u32(kf["node_feats"])]


def search_by_bow(kf1, second, second_is_frame, nnratio, check_orientation):
    a, b = _pk(kf1), _pk(second)
    n_out = len(b[0]) if second_is_frame else len(a[0])
    m = np.zeros(max(n_out, 1), np.int32)
    nm = lib().ygzr_search_by_bow(_p(a[0]), _p(a[1]), _p(a[2]), len(a[0]), _p(a[3]), _p(a[4]), _p(a[5]), len(a[3]),
                                  _p(b[0]), _p(b[1]), _p(b[2]), len(b[0]), _p(b[3]), _p(b[4]), _p(b[5]), len(b[3]),
                                  int(second_is_frame), cf(nnratio), int(check_orientation), _p(m))
    return nm, m[:n_out].copy()


def search_for_triangulation(kf1, kf2, F12, ex, ey, only_stereo, check_orientation):
    a, b = _pk(kf1, True), _pk(kf2, True)
    sf2, s2, F = f32(kf2["scale_factors"]), f32(kf2["level_sigma2"]), f32(F12).reshape(9)
    m12 = np.zeros(max(len(a[0]), 1), np.int32)
    nm = lib().ygzr_search_for_triangulation(
        _p(a[0]), _p(a[1]), _p(a[2]), _p(a[3]), len(a[0]), _p(a[4]), _p(a[5]), _p(a[6]), len(a[4]),
        _p(b[0]), _p(b[1]), _p(b[2]), _p(b[3]), len(b[0]), _p(b[4]), _p(b[5]), _p(b[6]), len(b[4]),
        _p(sf2), _p(s2), _p(F), cf(ex), cf(ey), int(only_stereo), int(check_orientation), _p(m12))
    return nm, m12[:len(a[0])].copy()


def stereo_match(kps_l, desc_l, kps_r, desc_r, pyr_l, pyr_r, scale_factors, inv_scale_factors, bf, b):
    """Frame::ComputeStereoMatches.  Returns (u_right, depth)."""
    kps_l, kps_r = np.ascontiguousarray(kps_l), np.ascontiguousarray(kps_r)
    desc_l, desc_r = u8(desc_l), u8(desc_r)
    lw = np.array([l.shape[1] for l in pyr_l], np.int32)
    lh = np.array([l.shape[0] for l in pyr_l], np.int32)
    off = np.zeros(len(pyr_l), np.int64)
    off[1:] = np.cumsum(lw[:-1].astype(np.int64) * lh[:-1])
    pl = np.concatenate([u8(l).ravel() for l in pyr_l])
    pr = np.concatenate([u8(l).ravel() for l in pyr_r])
    n = len(kps_l)
    ur, dp = np.zeros(max(n, 1), np.float32), np.zeros(max(n, 1), np.float32)
    sf, isf = f32(scale_factors), f32(inv_scale_factors)
    lib().ygzr_stereo_match(_p(kps_l), _p(desc_l), n, _p(kps_r), _p(desc_r), len(kps_r), _p(pl), _p(pr), _p(off), _p(lw), _p(lh),
                            len(pyr_l), _p(sf), _p(isf), cf(bf), cf(b), _p(ur), _p(dp))
    return ur[:n].copy(), dp[:n].copy()


def distinctive_descriptors(obs_desc, obs_start):
    """MapPoint::ComputeDistinctiveDescriptors per point.  Returns (changed, desc)."""
    obs_desc = u8(obs_desc).reshape(-1, 32)
    obs_start = i32(obs_start)
    n = len(obs_start) - 1
    out = np.zeros((max(n, 1), 32), np.uint8)
    ch = np.zeros(max(n, 1), np.uint8)
    lib().ygzr_distinctive_descriptors(_p(obs_desc), _p(obs_start), n, _p(out), _p(ch))
    return ch[:n].copy(), out[:n].copy()


class DirectKeyFrame(C.Structure):
    _fields_ = [("levels", C.c_void_p * 8), ("w", C.c_int32 * 8), ("h", C.c_int32 * 8), ("pitch", C.c_int32 * 8),
                ("pose_ref", C.c_float * 7), ("t_cur_ref", C.c_float * 7),
                ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float)]


def find_direct_projection(kfs, cur_levels, cur_intr, cur_pose, scale, inv_scale, inv_sigma2, kf_index, px_ref, octave, world_pos,
                           px_curr):
    """ORBmatcher::FindDirectProjection per (MapPoint, keyframe) pair.  kfs: list of dicts {levels, pose_ref, fx, fy, cx, cy};
    cur_pose = curr->mTcw (quaternion x, y, z, w + translation).  Returns (px, level, ok, t_cur_ref per keyframe): t_cur_ref is
    the TCR = curr->mTcw * pose_ref.inverse() the reference computed (src/ORBmatcher.cc:1582)."""
    keep = []
    arr = (DirectKeyFrame * max(len(kfs), 1))()
    for a, k in zip(arr, kfs):
        imgs = [u8(x) for x in k["levels"]]
        keep.append(imgs)
        for l, im in enumerate(imgs):
            a.levels[l] = im.ctypes.data; a.w[l] = im.shape[1]; a.h[l] = im.shape[0]; a.pitch[l] = im.shape[1]
        a.pose_ref[:] = [float(x) for x in k["pose_ref"]]
        a.fx, a.fy, a.cx, a.cy = (float(k[n]) for n in ("fx", "fy", "cx", "cy"))
    cur = [u8(x) for x in cur_levels]
    ptrs = (C.c_void_p * len(cur))(*[x.ctypes.data for x in cur])
    cw = np.array([x.shape[1] for x in cur], np.int32); ch = np.array([x.shape[0] for x in cur], np.int32)
    kf_index, octave = i32(kf_index), i32(octave)
    px_ref, world_pos = f32(px_ref), f32(world_pos)
    out = np.array(px_curr, np.float32).reshape(-1, 2).copy()
    n = len(kf_index)
    level, ok = np.zeros(n, np.int32), np.zeros(n, np.uint8)
    a = [f32(cur_intr), f32(cur_pose), f32(scale), f32(inv_scale), f32(inv_sigma2)]
    lib().ygzr_find_direct_projection(arr, len(kfs), ptrs, _p(cw), _p(ch), _p(cw), len(cur), *[_p(x) for x in a], n, _p(kf_index),
                                      _p(px_ref), _p(octave), _p(world_pos), _p(out), _p(level), _p(ok))
    tcr = [np.array(list(arr[k].t_cur_ref), np.float32) for k in range(len(kfs))]
    return out, level, ok, tcr


def search_by_projection_batch(n_kp, n_mp, kps, desc, bounds, scale_factors, mp, th, check_level, nnratio, threads=1, repeat=1):
    """The reference's SearchByProjection over a batch at fixed strides (layout of oracle_lib.search_by_projection_batch).
    Returns (kp_to_mp, n_matches, seconds): seconds = the timed region inside the library (grid assignment + search)."""
    F, ks = kps.shape
    ms = mp["proj_x"].shape[1]
    inv_w, inv_h = grid_inv(*bounds)
    sf = f32(scale_factors)
    kp_to_mp = np.full((F, ks), -1, np.int32)
    nm = np.zeros(F, np.int32)
    fn = lib().ygzr_search_by_projection_batch
    fn.restype = C.c_double
    pxr = mp.get("proj_xr")
    sec = fn(F, ks, ms, _p(i32(n_kp)), _p(i32(n_mp)), _p(np.ascontiguousarray(kps)), _p(u8(desc)), cf(bounds[0]), cf(bounds[1]),
             cf(inv_w), cf(inv_h), _p(sf), len(sf), _p(f32(mp["proj_x"])), _p(f32(mp["proj_y"])), _p(f32(pxr)) if pxr is not None else None,
             _p(f32(mp["view_cos"])), _p(i32(mp["level"])), _p(u8(mp["in_view"])), _p(u8(mp["has_obs"])), _p(u8(mp["desc"])),
             cf(th), int(check_level), cf(nnratio), int(threads), int(repeat), _p(kp_to_mp), _p(nm))
    return kp_to_mp, nm, float(sec)


class RPyramid(C.Structure):   # = YgzrPyramid (oracle/ref_match_api.cpp)
    _fields_ = [("levels", C.c_void_p * 8), ("w", C.c_int32 * 8), ("h", C.c_int32 * 8), ("pitch", C.c_int32 * 8)]


def _rpyr(levels):
    imgs = [np.ascontiguousarray(x, np.uint8) for x in levels]
    P = RPyramid()
    for l, im in enumerate(imgs):
        P.levels[l] = im.ctypes.data; P.w[l] = im.shape[1]; P.h[l] = im.shape[0]; P.pitch[l] = im.shape[1]
    return P, imgs


def sparse_align_residuals(ref_levels, cur_levels, level, inv_scale, focal, cur_intr, ref_pose, t_cur_ref, kp_xy, valid, world_pos):
    """precomputeReferencePatches + computeResiduals(linearize) of the reference's SparseImgAlign at one level.
    Returns H, Jres, chi2 (mean, as the reference returns it), n_meas, visible, ref_patch_cache, jacobian_cache."""
    R, k1 = _rpyr(ref_levels); Cu, k2 = _rpyr(cur_levels)
    kp_xy, valid, world_pos = f32(kp_xy), u8(valid), f32(world_pos)
    n = len(valid)
    H, J, chi2, nm = np.zeros((6, 6), np.float32), np.zeros(6, np.float32), C.c_float(), C.c_int64()
    vis, patch, jac = np.zeros(n, np.uint8), np.zeros((n, 16), np.float32), np.zeros((n, 16, 6), np.float32)
    lib().ygzr_sparse_align_residuals(C.byref(R), C.byref(Cu), len(ref_levels), level, _p(f32(inv_scale)), cf(focal), _p(f32(cur_intr)),
                                      _p(f32(ref_pose)), _p(f32(t_cur_ref)), n, _p(kp_xy), _p(valid), _p(world_pos), _p(H), _p(J),
                                      C.byref(chi2), C.byref(nm), _p(vis), _p(patch), _p(jac))
    return H, J, chi2.value, nm.value, vis, patch, jac


def sparse_align_run(ref_levels, cur_levels, max_level, min_level, inv_scale, focal, cur_intr, ref_pose, cur_pose, kp_xy, valid,
                     world_pos):
    """SparseImgAlign::run(ref, cur, TCR).  Returns (t_cur_ref0 = cur->mTcw * ref->mTcw.inverse() as composed by the reference,
    TCR, return value, [chi2_, n_meas_])."""
    R, k1 = _rpyr(ref_levels); Cu, k2 = _rpyr(cur_levels)
    kp_xy, valid, world_pos = f32(kp_xy), u8(valid), f32(world_pos)
    T0, T, stats = np.zeros(7, np.float32), np.zeros(7, np.float32), np.zeros(2, np.float32)
    lib().ygzr_sparse_align_run.restype = C.c_int64
    ret = lib().ygzr_sparse_align_run(C.byref(R), C.byref(Cu), len(ref_levels), max_level, min_level, _p(f32(inv_scale)), cf(focal),
                                      _p(f32(cur_intr)), _p(f32(ref_pose)), _p(f32(cur_pose)), len(valid), _p(kp_xy), _p(valid),
                                      _p(world_pos), _p(T0), _p(T), _p(stats))
    return T0, T, ret, stats
