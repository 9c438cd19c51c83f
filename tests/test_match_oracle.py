"""CPU checks of the matcher oracle (oracle/match_oracle.cpp): 'parity unpinned' by the reference,
so each function is compared with an independent brute-force re-derivation in numpy / pure Python
written from the reference's rules, on small seeded cases."""
import math

import numpy as np
import pytest

import oracle_lib as O
import synth


def _hamming_matrix(a, b):
    x = np.bitwise_xor(a[:, None, :], b[None, :, :])
    return np.unpackbits(x, axis=2).sum(axis=2).astype(np.int32)


def _features(w=320, h=240, seed=3, nf=400):
    img = synth.textured_frame(w, h, seed)
    return img, O.extract(O.params(nf), img)


def test_descriptor_distance():
    rng = np.random.default_rng(0)
    a = rng.integers(0, 256, size=(50, 32), dtype=np.uint8)
    b = rng.integers(0, 256, size=(50, 32), dtype=np.uint8)
    d = _hamming_matrix(a, b)
    for i in range(50):
        assert O.descriptor_distance(a[i], b[i]) == d[i, i]
    assert O.descriptor_distance(a[0], a[0]) == 0
    assert O.descriptor_distance(np.zeros(32, np.uint8), np.full(32, 255, np.uint8)) == 256


@pytest.mark.parametrize("nq,nt,ratio,th", [(200, 300, 0.6, 50), (64, 1, 0.9, 100), (100, 257, 0.75, 256), (5, 0, 0.6, 50)])
def test_knn2_vs_numpy(nq, nt, ratio, th):
    q, t = synth.knn_case(nq, max(nt, 1), 5)
    t = t[:nt]
    idx, d1, d2, match = O.hamming_knn2(q, t, ratio, th, threads=2)
    if nt == 0:
        assert (idx == -1).all() and (d1 == 256).all() and (match == -1).all()
        return
    D = _hamming_matrix(q, t)
    bi = D.argmin(axis=1)                       # first index of the minimum
    assert np.array_equal(idx, bi) and np.array_equal(d1, D[np.arange(nq), bi])
    D2 = D.copy(); D2[np.arange(nq), bi] = 257
    exp2 = np.minimum(D2.min(axis=1), 256) if nt > 1 else np.full(nq, 256)
    assert np.array_equal(d2, exp2)
    ok = (d1 <= th) & (d1.astype(np.float32) < np.float32(ratio) * d2.astype(np.float32))
    assert np.array_equal(match, np.where(ok, bi, -1))


def _py_search_by_projection(kps, desc, u_right, claimed, bounds, sf, mp, th, check_level, ratio):
    """Independent pure-Python restatement of src/ORBmatcher.cc:44-126 + src/Frame.cc:424-493."""
    f = np.float32
    min_x, min_y, max_x, max_y = [f(v) for v in bounds]
    inv_w, inv_h = f(64) / (max_x - min_x), f(48) / (max_y - min_y)
    rnd = lambda v: int(math.floor(abs(v) + 0.5)) * (1 if v >= 0 else -1)
    grid = {}
    for i in range(len(kps)):
        px, py = rnd(float((kps["x"][i] - min_x) * inv_w)), rnd(float((kps["y"][i] - min_y) * inv_h))
        if 0 <= px < 64 and 0 <= py < 48:
            grid.setdefault((px, py), []).append(i)
    D = _hamming_matrix(mp["desc"], desc)
    blocked = claimed.astype(bool).copy() if claimed is not None else np.zeros(len(kps), bool)
    kp_to_mp = np.full(len(kps), -1, np.int32)
    best_of = np.full(len(mp["proj_x"]), -1, np.int32)
    n = 0
    for i in range(len(mp["proj_x"])):
        if not mp["in_view"][i]:
            continue
        lvl = int(mp["level"][i])
        r = f(2.5) if float(mp["view_cos"][i]) > 0.998 else f(4.0)
        if th != 1.0:
            r = f(r * f(th))
        rad = f(r * sf[lvl])
        x, y = f(mp["proj_x"][i]), f(mp["proj_y"][i])
        cx0 = max(0, math.floor(float(f(f(x - min_x - rad) * inv_w)))); cx1 = min(63, math.ceil(float(f(f(x - min_x + rad) * inv_w))))
        cy0 = max(0, math.floor(float(f(f(y - min_y - rad) * inv_h)))); cy1 = min(47, math.ceil(float(f(f(y - min_y + rad) * inv_h))))
        if cx0 >= 64 or cx1 < 0 or cy0 >= 48 or cy1 < 0:
            continue
        b1, b2, l1, l2, bi = 256, 256, -1, -1, -1
        found = False
        for ix in range(cx0, cx1 + 1):
            for iy in range(cy0, cy1 + 1):
                for idx in grid.get((ix, iy), []):
                    o = int(kps["octave"][idx])
                    if check_level and (o < lvl - 1 or o > lvl):
                        continue
                    if not (abs(f(kps["x"][idx] - x)) < rad and abs(f(kps["y"][idx] - y)) < rad):
                        continue
                    found = True
                    if blocked[idx]:
                        continue
                    if u_right is not None and u_right[idx] > 0 and abs(f(mp["proj_xr"][i] - u_right[idx])) > rad:
                        continue
                    d = int(D[i, idx])
                    if d < b1:
                        b2, b1, l2, l1, bi = b1, d, l1, o, idx
                    elif d < b2:
                        l2, b2 = o, d
        if not found:
            continue
        if b1 <= 100:
            if l1 == l2 and f(b1) > f(ratio) * f(b2):
                continue
            kp_to_mp[bi] = i; blocked[bi] = bool(mp["has_obs"][i]); best_of[i] = bi; n += 1
    return n, kp_to_mp, best_of


@pytest.mark.parametrize("th,check_level,stereo,seed", [(1.0, False, False, 1), (5.0, False, True, 2), (3.0, True, False, 3),
                                                         (15.0, False, False, 4)])
def test_search_by_projection_vs_python(th, check_level, stereo, seed):
    img, (kps, desc) = _features(seed=seed)
    sf, _, _ = synth.scale_tables()
    mp, u_right, claimed = synth.projection_case(kps, desc, 600, seed, stereo=stereo)
    bounds = (0.0, 0.0, float(img.shape[1]), float(img.shape[0]))
    a = O.search_by_projection(kps, desc, u_right, claimed, bounds, sf, mp, th, check_level, 0.8)
    b = _py_search_by_projection(kps, desc, u_right, claimed, bounds, sf, mp, th, check_level, 0.8)
    assert a[0] == b[0] and a[0] > 50
    assert np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])


def _py_search_by_projection_last(kps, desc, u_right, claimed, bounds, sf, last, th, mono, check_level, check_ori):
    """src/ORBmatcher.cc:1218-1350 re-derived in float32 numpy scalars (dot products as a0 + (a1 + a2))."""
    f = np.float32
    min_x, min_y, max_x, max_y = [f(b) for b in bounds]
    inv_w, inv_h = f(64) / (max_x - min_x), f(48) / (max_y - min_y)
    rnd = lambda v: int(math.floor(abs(v) + 0.5)) * (1 if v >= 0 else -1)
    grid = {}
    for i in range(len(kps)):
        px, py = rnd(float((kps["x"][i] - min_x) * inv_w)), rnd(float((kps["y"][i] - min_y) * inv_h))
        if 0 <= px < 64 and 0 <= py < 48:
            grid.setdefault((px, py), []).append(i)
    P = last["pose"].astype(np.float32)
    Rcw, tcw, Rlw, tlw = P[0:9].reshape(3, 3), P[9:12], P[12:21].reshape(3, 3), P[21:24]
    fx, fy, cx, cy, mbf, mb = P[24:30]
    dot3 = lambda a0, a1, a2: f(a0 + f(a1 + a2))
    twc = [f(-dot3(f(Rcw[0, r] * tcw[0]), f(Rcw[1, r] * tcw[1]), f(Rcw[2, r] * tcw[2]))) for r in range(3)]
    tlcz = f(dot3(f(Rlw[2, 0] * twc[0]), f(Rlw[2, 1] * twc[1]), f(Rlw[2, 2] * twc[2])) + tlw[2])
    fwd, bwd = (tlcz > mb and not mono), (-tlcz > mb and not mono)
    D = _hamming_matrix(last["desc"], desc)
    blocked = claimed.astype(bool).copy() if claimed is not None else np.zeros(len(kps), bool)
    kp_to_pt = np.full(len(kps), -1, np.int32)
    best_of = np.full(len(last["valid"]), -1, np.int32)
    hist = [[] for _ in range(30)]
    n = 0
    for i in range(len(last["valid"])):
        if not last["valid"][i]:
            continue
        X = last["world_pos"][i].astype(np.float32)
        c = [f(dot3(f(Rcw[r, 0] * X[0]), f(Rcw[r, 1] * X[1]), f(Rcw[r, 2] * X[2])) + tcw[r]) for r in range(3)]
        with np.errstate(divide="ignore"):
            invz = f(np.float64(1.0) / np.float64(c[2]))
        if invz < 0:
            continue
        u, v = f(f(f(fx * c[0]) * invz) + cx), f(f(f(fy * c[1]) * invz) + cy)
        if not (np.isfinite(u) and np.isfinite(v)) or u < min_x or u > max_x or v < min_y or v > max_y:
            continue
        o = int(last["octave"][i])
        rad = f(f(th) * sf[o])
        if not check_level:
            lo, hi = -1, -1
        elif fwd:
            lo, hi = o, -1
        elif bwd:
            lo, hi = 0, o
        else:
            lo, hi = o - 1, o + 1
        chk = lo > 0 or hi >= 0
        cx0 = max(0, math.floor(float(f(f(u - min_x - rad) * inv_w)))); cx1 = min(63, math.ceil(float(f(f(u - min_x + rad) * inv_w))))
        cy0 = max(0, math.floor(float(f(f(v - min_y - rad) * inv_h)))); cy1 = min(47, math.ceil(float(f(f(v - min_y + rad) * inv_h))))
        if cx0 >= 64 or cx1 < 0 or cy0 >= 48 or cy1 < 0:
            continue
        b1, bi = 256, -1
        for ix in range(cx0, cx1 + 1):
            for iy in range(cy0, cy1 + 1):
                for idx in grid.get((ix, iy), []):
                    ko = int(kps["octave"][idx])
                    if chk and (ko < lo or (hi >= 0 and ko > hi)):
                        continue
                    if not (abs(f(kps["x"][idx] - u)) < rad and abs(f(kps["y"][idx] - v)) < rad):
                        continue
                    if blocked[idx]:
                        continue
                    if u_right is not None and u_right[idx] > 0 and abs(f(f(u - f(mbf * invz)) - u_right[idx])) > rad:
                        continue
                    d = int(D[i, idx])
                    if d < b1:
                        b1, bi = d, idx
        if b1 <= 100:
            kp_to_pt[bi] = i; blocked[bi] = bool(last["has_obs"][i]); best_of[i] = bi; n += 1
            if check_ori:
                rot = f(last["angle"][i] - kps["angle"][bi])
                if rot < 0:
                    rot = f(rot + f(360))
                b = rnd(float(f(rot * f(f(1.0) / f(30)))))
                hist[0 if b == 30 else b].append(bi)
    if check_ori:
        m1 = m2 = m3 = 0; i1 = i2 = i3 = -1
        for i in range(30):
            sz = len(hist[i])
            if sz > m1:
                m3, m2, m1, i3, i2, i1 = m2, m1, sz, i2, i1, i
            elif sz > m2:
                m3, m2, i3, i2 = m2, sz, i2, i
            elif sz > m3:
                m3, i3 = sz, i
        if f(m2) < f(0.1) * f(m1):
            i2 = i3 = -1
        elif f(m3) < f(0.1) * f(m1):
            i3 = -1
        for i in range(30):
            if i not in (i1, i2, i3):
                for idx in hist[i]:
                    kp_to_pt[idx] = -1; n -= 1
    return n, kp_to_pt, best_of


@pytest.mark.parametrize("th,mono,check_level,check_ori,stereo,motion,seed", [
    (15.0, True, False, True, False, (0.02, -0.01, 0.3), 1), (7.0, False, True, True, True, (0.0, 0.0, 0.5), 2),
    (7.0, False, True, False, True, (0.0, 0.0, -0.5), 3), (15.0, False, True, True, False, (0.05, 0.0, 0.01), 4)])
def test_search_by_projection_last_vs_python(th, mono, check_level, check_ori, stereo, motion, seed):
    img, (kps, desc) = _features(seed=seed)
    sf, _, _ = synth.scale_tables()
    last, u_right, claimed = synth.lastframe_case(kps, desc, 500, seed, motion=motion, stereo=stereo)
    bounds = (0.0, 0.0, float(img.shape[1]), float(img.shape[0]))
    a = O.search_by_projection_last(kps, desc, u_right, claimed, bounds, sf, last, th, mono, check_level, check_ori)
    b = _py_search_by_projection_last(kps, desc, u_right, claimed, bounds, sf, last, th, mono, check_level, check_ori)
    assert a[0] == b[0] and a[0] > 30
    assert np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
    if check_ori:   # the rotation vote must have removed something, and never more than it added
        assert (a[2] >= 0).sum() > a[0]


def _py_search_by_projection_keyframe(kps, desc, claimed, bounds, sf, pts, th, variant, orb_dist, check_ori, u_right=None,
                                      inv_sigma2=None):
    """src/ORBmatcher.cc:1352-1469 (variant 0, Eigen order), :265-373 (1), Fuse :748-886 (2, Eigen), Fuse with Sim3 :888-1004
    (3) and one direction of SearchBySim3 :1043-1110 (4) re-derived; the cv::Mat variants call the installed OpenCV
    (cv2.gemm, cv2.norm) for those expressions, which pins the oracle's restatement of them."""
    import cv2
    f = np.float32
    variant = int(variant)
    eigen, claims = variant in (0, 2), variant <= 1
    min_x, min_y, max_x, max_y = [f(b) for b in bounds]
    inv_w, inv_h = f(64) / (max_x - min_x), f(48) / (max_y - min_y)
    rnd = lambda v: int(math.floor(abs(v) + 0.5)) * (1 if v >= 0 else -1)
    grid = {}
    for i in range(len(kps)):
        px, py = rnd(float((kps["x"][i] - min_x) * inv_w)), rnd(float((kps["y"][i] - min_y) * inv_h))
        if 0 <= px < 64 and 0 <= py < 48:
            grid.setdefault((px, py), []).append(i)
    P = pts["pose"].astype(np.float32)
    Rcw, tcw = P[0:9].reshape(3, 3), P[9:12]
    fx, fy, cx, cy = P[24:28]
    dot3 = lambda a0, a1, a2: f(a0 + f(a1 + a2))
    Ow = P[12:15] if variant != 0 else np.array([f(-dot3(f(Rcw[0, r] * tcw[0]), f(Rcw[1, r] * tcw[1]), f(Rcw[2, r] * tcw[2]))) for r in range(3)], f)
    D = _hamming_matrix(pts["desc"], desc)
    blocked = claimed.astype(bool).copy() if (claimed is not None and claims) else np.zeros(len(kps), bool)
    kp_to_pt = np.full(len(kps), -1, np.int32)
    best_of = np.full(len(pts["valid"]), -1, np.int32)
    hist = [[] for _ in range(30)]
    n = 0
    log_sf, nlev = f(pts["log_scale_factor"]), int(pts["n_scale_levels"])
    for i in range(len(pts["valid"])):
        if not pts["valid"][i]:
            continue
        X = pts["world_pos"][i].astype(np.float32)
        with np.errstate(divide="ignore", invalid="ignore"):
            if eigen:
                c = [f(dot3(f(Rcw[r, 0] * X[0]), f(Rcw[r, 1] * X[1]), f(Rcw[r, 2] * X[2])) + tcw[r]) for r in range(3)]
            else:
                c = cv2.gemm(Rcw, X.reshape(3, 1), 1.0, tcw.reshape(3, 1), 1.0)[:, 0]
            if variant == 4:
                c = cv2.gemm(P[12:21].reshape(3, 3), c.reshape(3, 1), 1.0, P[21:24].reshape(3, 1), 1.0)[:, 0]
            if variant == 0:
                invz = f(np.float64(1.0) / np.float64(c[2]))
                u, v = f(f(f(fx * c[0]) * invz) + cx), f(f(f(fy * c[1]) * invz) + cy)
                if not (np.isfinite(u) and np.isfinite(v)) or u < min_x or u > max_x or v < min_y or v > max_y:
                    continue
            else:
                if c[2] < 0.0:
                    continue
                invz = f(1) / f(c[2]) if variant in (1, 2) else f(np.float64(1.0) / np.float64(c[2]))
                u, v = f(f(fx * f(c[0] * invz)) + cx), f(f(fy * f(c[1] * invz)) + cy)
                if not (u >= min_x and u < max_x and v >= min_y and v < max_y):
                    continue
            ur = f(u - f(P[28] * invz))
            PO = np.array(c, f) if variant == 4 else X - Ow
            if eigen:
                dist = f(np.sqrt(dot3(f(PO[0] * PO[0]), f(PO[1] * PO[1]), f(PO[2] * PO[2]))))
            else:
                dist = f(cv2.norm(PO.reshape(3, 1)))
        if dist < f(f(0.8) * pts["min_distance"][i]) or dist > f(f(1.2) * pts["max_distance"][i]):
            continue
        if variant in (1, 2, 3):
            Pn = pts["normal"][i]
            if variant == 2:
                d = np.float64(dot3(f(PO[0] * Pn[0]), f(PO[1] * Pn[1]), f(PO[2] * Pn[2])))
            else:
                d = np.float64(0)
                for r in range(3):
                    d = d + np.float64(PO[r]) * np.float64(Pn[r])
            if d < 0.5 * np.float64(dist):
                continue
        lvl = int(math.ceil(float(f(f(np.log(f(pts["max_distance"][i] / dist))) / log_sf))))
        lvl = min(max(lvl, 0), nlev - 1)
        rad = f(f(th) * sf[lvl])
        lo, hi = lvl - 1, (lvl + 1 if variant == 0 else lvl)
        cx0 = max(0, math.floor(float(f(f(u - min_x - rad) * inv_w)))); cx1 = min(63, math.ceil(float(f(f(u - min_x + rad) * inv_w))))
        cy0 = max(0, math.floor(float(f(f(v - min_y - rad) * inv_h)))); cy1 = min(47, math.ceil(float(f(f(v - min_y + rad) * inv_h))))
        if cx0 >= 64 or cx1 < 0 or cy0 >= 48 or cy1 < 0:
            continue
        b1, bi = 256, -1
        for ix in range(cx0, cx1 + 1):
            for iy in range(cy0, cy1 + 1):
                for idx in grid.get((ix, iy), []):
                    ko = int(kps["octave"][idx])
                    if ko < lo or ko > hi:
                        continue
                    if not (abs(f(kps["x"][idx] - u)) < rad and abs(f(kps["y"][idx] - v)) < rad):
                        continue
                    if blocked[idx]:
                        continue
                    if variant == 2:
                        ex, ey = f(u - kps["x"][idx]), f(v - kps["y"][idx])
                        e2 = f(f(ex * ex) + f(ey * ey))
                        if u_right is not None and u_right[idx] >= 0:
                            er = f(ur - u_right[idx])
                            if np.float64(f(f(e2 + f(er * er)) * inv_sigma2[ko])) > 7.8:
                                continue
                        elif np.float64(f(e2 * inv_sigma2[ko])) > 5.99:
                            continue
                    if int(D[i, idx]) < b1:
                        b1, bi = int(D[i, idx]), idx
        if b1 <= orb_dist:
            kp_to_pt[bi] = i; blocked[bi] = claims; best_of[i] = bi; n += 1
            if check_ori:
                rot = f(pts["angle"][i] - kps["angle"][bi])
                if rot < 0:
                    rot = f(rot + f(360))
                b = rnd(float(f(rot * f(f(1.0) / f(30)))))
                hist[0 if b == 30 else b].append(bi)
    if check_ori:
        m1 = m2 = m3 = 0; i1 = i2 = i3 = -1
        for i in range(30):
            sz = len(hist[i])
            if sz > m1:
                m3, m2, m1, i3, i2, i1 = m2, m1, sz, i2, i1, i
            elif sz > m2:
                m3, m2, i3, i2 = m2, sz, i2, i
            elif sz > m3:
                m3, i3 = sz, i
        if f(m2) < f(0.1) * f(m1):
            i2 = i3 = -1
        elif f(m3) < f(0.1) * f(m1):
            i3 = -1
        for i in range(30):
            if i not in (i1, i2, i3):
                for idx in hist[i]:
                    kp_to_pt[idx] = -1; n -= 1
    return n, kp_to_pt, best_of


@pytest.mark.parametrize("th,variant,orb_dist,check_ori,seed", [
    (10.0, 0, 100, True, 1), (3.0, 0, 64, True, 2), (10.0, 1, 50, False, 3), (4.0, 1, 50, False, 4), (3.0, 2, 50, False, 5),
    (4.0, 3, 50, False, 6), (7.5, 4, 100, False, 7)])
def test_search_by_projection_keyframe_vs_python(th, variant, orb_dist, check_ori, seed):
    img, (kps, desc) = _features(seed=seed)
    sf, _, sigma2 = synth.scale_tables()
    inv_sigma2 = (np.float32(1.0) / sigma2).astype(np.float32)
    pts, claimed = synth.keyframe_points_case(kps, desc, 500, seed, variant, cx=160.0, cy=120.0)
    u_right = pts["u_right"] if variant == 2 else None
    bounds = (0.0, 0.0, float(img.shape[1]), float(img.shape[0]))
    a = O.search_by_projection_keyframe(kps, desc, claimed, bounds, sf, pts, th, variant, orb_dist, check_ori, u_right, inv_sigma2)
    b = _py_search_by_projection_keyframe(kps, desc, claimed, bounds, sf, pts, th, variant, orb_dist, check_ori, u_right,
                                          inv_sigma2)
    assert a[0] == b[0] and a[0] > 30
    assert np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
    if check_ori:
        assert (a[2] >= 0).sum() > a[0]


def _py_search_for_initialization(k1, d1, k2, d2, bounds, prev, window, ratio, check_ori):
    """src/ORBmatcher.cc:375-478 re-derived."""
    f = np.float32
    min_x, min_y, max_x, max_y = [f(b) for b in bounds]
    inv_w, inv_h = f(64) / (max_x - min_x), f(48) / (max_y - min_y)
    rnd = lambda v: int(math.floor(abs(v) + 0.5)) * (1 if v >= 0 else -1)
    grid = {}
    for i in range(len(k2)):
        px, py = rnd(float((k2["x"][i] - min_x) * inv_w)), rnd(float((k2["y"][i] - min_y) * inv_h))
        if 0 <= px < 64 and 0 <= py < 48:
            grid.setdefault((px, py), []).append(i)
    D = _hamming_matrix(d1, d2)
    prev = prev.copy()
    m12 = np.full(len(k1), -1, np.int32); m21 = np.full(len(k2), -1, np.int32)
    md = np.full(len(k2), 2**31 - 1, np.int64)
    hist = [[] for _ in range(30)]
    n = 0
    r = f(window)
    for i1 in range(len(k1)):
        if k1["octave"][i1] > 0:
            continue
        x, y = prev[i1]
        cx0 = max(0, math.floor(float(f(f(x - min_x - r) * inv_w)))); cx1 = min(63, math.ceil(float(f(f(x - min_x + r) * inv_w))))
        cy0 = max(0, math.floor(float(f(f(y - min_y - r) * inv_h)))); cy1 = min(47, math.ceil(float(f(f(y - min_y + r) * inv_h))))
        if cx0 >= 64 or cx1 < 0 or cy0 >= 48 or cy1 < 0:
            continue
        b1 = b2 = 2**31 - 1; bi = -1
        for ix in range(cx0, cx1 + 1):
            for iy in range(cy0, cy1 + 1):
                for i2 in grid.get((ix, iy), []):
                    if k2["octave"][i2] != 0:
                        continue
                    if not (abs(f(k2["x"][i2] - x)) < r and abs(f(k2["y"][i2] - y)) < r):
                        continue
                    d = int(D[i1, i2])
                    if md[i2] <= d:
                        continue
                    if d < b1:
                        b2, b1, bi = b1, d, i2
                    elif d < b2:
                        b2 = d
        if b1 <= 50 and f(b1) < f(f(b2) * f(ratio)):
            if m21[bi] >= 0:
                m12[m21[bi]] = -1; n -= 1
            m12[i1] = bi; m21[bi] = i1; md[bi] = b1; n += 1
            if check_ori:
                rot = f(k1["angle"][i1] - k2["angle"][bi])
                if rot < 0:
                    rot = f(rot + f(360))
                b = rnd(float(f(rot * f(f(1.0) / f(30)))))
                hist[0 if b == 30 else b].append(i1)
    if check_ori:
        m1 = m2 = m3 = 0; j1 = j2 = j3 = -1
        for i in range(30):
            sz = len(hist[i])
            if sz > m1:
                m3, m2, m1, j3, j2, j1 = m2, m1, sz, j2, j1, i
            elif sz > m2:
                m3, m2, j3, j2 = m2, sz, j2, i
            elif sz > m3:
                m3, j3 = sz, i
        if f(m2) < f(0.1) * f(m1):
            j2 = j3 = -1
        elif f(m3) < f(0.1) * f(m1):
            j3 = -1
        for i in range(30):
            if i not in (j1, j2, j3):
                for idx1 in hist[i]:
                    if m12[idx1] >= 0:
                        m12[idx1] = -1; n -= 1
    for i1 in range(len(k1)):
        if m12[i1] >= 0:
            prev[i1] = (k2["x"][m12[i1]], k2["y"][m12[i1]])
    return n, m12, prev


@pytest.mark.parametrize("window,ratio,check_ori,seed", [(100, 0.9, True, 1), (30, 0.9, True, 2), (100, 0.7, False, 3)])
def test_search_for_initialization_vs_python(window, ratio, check_ori, seed):
    img, (kps, desc) = _features(seed=seed)
    k1, d1, k2, d2, prev = synth.initialization_case(kps, desc, seed, n_dup=120)
    bounds = (0.0, 0.0, float(img.shape[1]), float(img.shape[0]))
    a = O.search_for_initialization(k1, d1, k2, d2, bounds, prev, window, ratio, check_ori)
    b = _py_search_for_initialization(k1, d1, k2, d2, bounds, prev, window, ratio, check_ori)
    assert a[0] == b[0] and a[0] > 40 and a[0] == (a[1] >= 0).sum()
    assert np.array_equal(a[1], b[1]) and a[2].tobytes() == b[2].tobytes()
    assert len(set(a[1][a[1] >= 0].tolist())) == a[0]          # a keypoint of F2 has one owner


def test_search_by_projection_order_dependence():
    """Two MapPoints with the same descriptor on the same keypoint: the first claims it, the second must skip it."""
    img, (kps, desc) = _features(seed=9)
    sf, _, _ = synth.scale_tables()
    k = 10
    mp = {"proj_x": np.array([kps["x"][k]] * 2, np.float32), "proj_y": np.array([kps["y"][k]] * 2, np.float32),
          "proj_xr": np.zeros(2, np.float32), "view_cos": np.array([0.95, 0.95], np.float32),
          "level": np.array([kps["octave"][k]] * 2, np.int32), "in_view": np.ones(2, np.uint8),
          "has_obs": np.array([1, 1], np.uint8), "desc": np.stack([desc[k], desc[k]])}
    nm, kp_to_mp, best = O.search_by_projection(kps, desc, None, None, (0, 0, img.shape[1], img.shape[0]), sf, mp, 1.0, False, 0.8)
    assert best[0] == k and kp_to_mp[k] == 0 and best[1] != k
    mp["has_obs"] = np.array([0, 1], np.uint8)          # first has no observations: the second overwrites it
    nm, kp_to_mp, best = O.search_by_projection(kps, desc, None, None, (0, 0, img.shape[1], img.shape[0]), sf, mp, 1.0, False, 0.8)
    assert best[0] == k and best[1] == k and kp_to_mp[k] == 1 and nm == 2


def _py_triangulation(kf1, kf2, F12, ex, ey, only_stereo, check_ori):
    f = np.float32
    D = _hamming_matrix(kf1["desc"], kf2["desc"])
    n1 = len(kf1["kps"])
    m12 = np.full(n1, -1, np.int32)
    nodes2 = {int(nid): k for k, nid in enumerate(kf2["node_ids"])}
    F = F12.astype(np.float32)
    bins = {}
    for a, nid in enumerate(kf1["node_ids"]):
        if int(nid) not in nodes2:
            continue
        b = nodes2[int(nid)]
        for idx1 in kf1["node_feats"][kf1["node_start"][a]:kf1["node_start"][a + 1]]:
            idx1 = int(idx1)
            if kf1["has_mp"][idx1]:
                continue
            s1 = kf1["u_right"][idx1] >= 0
            if only_stereo and not s1:
                continue
            x1, y1 = f(kf1["kps"]["x"][idx1]), f(kf1["kps"]["y"][idx1])
            best, bi = 50, -1
            for idx2 in kf2["node_feats"][kf2["node_start"][b]:kf2["node_start"][b + 1]]:
                idx2 = int(idx2)
                if kf2["has_mp"][idx2]:
                    continue
                s2 = kf2["u_right"][idx2] >= 0
                if only_stereo and not s2:
                    continue
                d = int(D[idx1, idx2])
                if d > 50 or d > best:
                    continue
                x2, y2, o2 = f(kf2["kps"]["x"][idx2]), f(kf2["kps"]["y"][idx2]), int(kf2["kps"]["octave"][idx2])
                if not s1 and not s2:
                    dx, dy = f(f(ex) - x2), f(f(ey) - y2)
                    if f(f(dx * dx) + f(dy * dy)) < f(f(100) * kf2["scale_factors"][o2]):
                        continue
                la = f(f(f(x1 * F[0, 0]) + f(y1 * F[1, 0])) + F[2, 0])
                lb = f(f(f(x1 * F[0, 1]) + f(y1 * F[1, 1])) + F[2, 1])
                lc = f(f(f(x1 * F[0, 2]) + f(y1 * F[1, 2])) + F[2, 2])
                num = f(f(f(la * x2) + f(lb * y2)) + lc)
                den = f(f(la * la) + f(lb * lb))
                if den == 0:
                    continue
                dsqr = f(f(num * num) / den)
                if float(dsqr) < 3.84 * float(kf2["level_sigma2"][o2]):
                    best, bi = d, idx2
            if bi >= 0:
                m12[idx1] = bi
                rot = f(kf1["kps"]["angle"][idx1] - kf2["kps"]["angle"][bi])
                if rot < 0:
                    rot = f(rot + f(360))
                v = float(f(rot * f(f(1.0) / f(30))))
                bn = int(math.floor(abs(v) + 0.5))
                bins.setdefault(0 if bn == 30 else bn, []).append(idx1)
    if check_ori:
        sizes = [len(bins.get(i, [])) for i in range(30)]
        m1 = m2 = m3 = 0; i1 = i2 = i3 = -1
        for i, s in enumerate(sizes):
            if s > m1:
                m3, m2, m1, i3, i2, i1 = m2, m1, s, i2, i1, i
            elif s > m2:
                m3, m2, i3, i2 = m2, s, i2, i
            elif s > m3:
                m3, i3 = s, i
        if m2 < f(0.1) * f(m1):
            i2 = i3 = -1
        elif m3 < f(0.1) * f(m1):
            i3 = -1
        for i in range(30):
            if i not in (i1, i2, i3):
                for idx1 in bins.get(i, []):
                    m12[idx1] = -1
    return int((m12 >= 0).sum()), m12


@pytest.mark.parametrize("only_stereo,check_ori,seed", [(False, False, 1), (False, True, 2), (True, True, 3)])
def test_triangulation_vs_python(only_stereo, check_ori, seed):
    _, (kps, desc) = _features(seed=seed + 10)
    kf1, kf2, F12, ex, ey = synth.triangulation_case(kps, desc, seed)
    a = O.search_for_triangulation(kf1, kf2, F12, ex, ey, only_stereo, check_ori)
    b = _py_triangulation(kf1, kf2, F12, ex, ey, only_stereo, check_ori)
    assert a[0] == b[0] and np.array_equal(a[1], b[1])
    assert a[0] > (5 if only_stereo else 30)


def _py_search_by_bow(kf1, kf2, frame2, ratio, check_ori):
    """src/ORBmatcher.cc:155-263 / :480-596 re-derived with dict FeatureVectors."""
    f = np.float32
    fv1 = {int(n): kf1["node_feats"][kf1["node_start"][k]:kf1["node_start"][k + 1]] for k, n in enumerate(kf1["node_ids"])}
    fv2 = {int(n): kf2["node_feats"][kf2["node_start"][k]:kf2["node_start"][k + 1]] for k, n in enumerate(kf2["node_ids"])}
    D = _hamming_matrix(kf1["desc"], kf2["desc"])
    n_out = len(kf2["kps"]) if frame2 else len(kf1["kps"])
    matches = np.full(n_out, -1, np.int32)
    taken = np.zeros(len(kf2["kps"]), bool)
    hist = [[] for _ in range(30)]
    rnd = lambda v: int(math.floor(abs(v) + 0.5)) * (1 if v >= 0 else -1)
    n = 0
    for node in sorted(set(fv1) & set(fv2)):
        for i1 in fv1[node]:
            if not kf1["has_mp"][i1]:
                continue
            b1, b2, bi = 256, 256, -1
            for i2 in fv2[node]:
                if taken[i2] or (not frame2 and not kf2["has_mp"][i2]):
                    continue
                d = int(D[i1, i2])
                if d < b1:
                    b2, b1, bi = b1, d, int(i2)
                elif d < b2:
                    b2 = d
            near = b1 <= 50 if frame2 else b1 < 50
            if near and f(b1) < f(ratio) * f(b2):
                taken[bi] = True
                slot = bi if frame2 else int(i1)
                matches[slot] = int(i1) if frame2 else bi
                n += 1
                if check_ori:
                    rot = f(kf1["kps"]["angle"][i1] - kf2["kps"]["angle"][bi])
                    if rot < 0:
                        rot = f(rot + f(360))
                    b = rnd(float(f(rot * f(f(1.0) / f(30)))))
                    hist[0 if b == 30 else b].append(slot)
    if check_ori:
        m1 = m2 = m3 = 0; i1 = i2 = i3 = -1
        for i in range(30):
            sz = len(hist[i])
            if sz > m1:
                m3, m2, m1, i3, i2, i1 = m2, m1, sz, i2, i1, i
            elif sz > m2:
                m3, m2, i3, i2 = m2, sz, i2, i
            elif sz > m3:
                m3, i3 = sz, i
        if f(m2) < f(0.1) * f(m1):
            i2 = i3 = -1
        elif f(m3) < f(0.1) * f(m1):
            i3 = -1
        for i in range(30):
            if i not in (i1, i2, i3):
                for slot in hist[i]:
                    matches[slot] = -1; n -= 1
    return n, matches


@pytest.mark.parametrize("frame2,check_ori,ratio,seed", [(True, True, 0.7, 1), (True, False, 0.9, 2), (False, True, 0.75, 3),
                                                        (False, False, 0.6, 4)])
def test_search_by_bow_vs_python(frame2, check_ori, ratio, seed):
    img, (kps, desc) = _features(seed=seed)
    kf1, kf2, _, _, _ = synth.triangulation_case(kps, desc, seed, n_nodes=40)
    kf1["has_mp"] = 1 - kf1["has_mp"]; kf2["has_mp"] = 1 - kf2["has_mp"]          # most keypoints carry a MapPoint
    a = O.search_by_bow(kf1, kf2, frame2, ratio, check_ori)
    b = _py_search_by_bow(kf1, kf2, frame2, ratio, check_ori)
    assert a[0] == b[0] and a[0] > 20 and np.array_equal(a[1], b[1])


def test_stereo_match_recovers_disparity():
    w, h = 480, 200
    left, right = synth.stereo_pair(w, h, 5)
    p = O.params(600)
    kl, dl = O.extract(p, left)
    kr, dr = O.extract(p, right)
    sf, isf, _ = synth.scale_tables()
    ur, depth, sad = O.stereo_match(kl, dl, kr, dr, O.pyramid(p, left), O.pyramid(p, right), sf, isf, bf=40.0, b=0.5)
    ok = ur >= 0
    assert ok.sum() > 100
    disp_true = 2 + 38 * (0.5 + 0.5 * np.sin(np.linspace(0, 3.0, h)))[np.clip(kl["y"][ok].astype(int), 0, h - 1)]
    err = np.abs((kl["x"][ok] - ur[ok]) - disp_true)
    assert np.median(err) < 1.0                         # sub-pixel refinement lands on the synthetic disparity
    assert np.allclose(depth[ok], np.float32(40.0) / (kl["x"][ok] - ur[ok]), rtol=1e-6)
    assert ((ur < 0) == (depth < 0)).all()
    # outlier cut: every surviving SAD is below 1.5*1.4*median of the accepted ones (src/Frame.cc:670-681)
    acc = sad[sad >= 0]
    med = np.sort(acc)[len(acc) // 2]
    assert (sad[ok] < np.float32(1.5) * np.float32(1.4) * np.float32(med)).all()
