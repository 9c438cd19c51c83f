"""Seeded synthetic inputs shared by the tests, bench.py and smoke().

Frames: uniform uint8 noise -> cv2/numpy Gaussian (sigma 2) -> min-max normalised to 0..255: dense
corners on every pyramid level.  `contrast` < 1 squeezes the grey range around 128 so that the
minThFAST fallback of the per-cell FAST (reference src/ORBextractor.cc:764-770) is exercised.
Pure numpy on purpose (no cv2) so the generator is identical wherever it runs.
"""
import numpy as np


def _gauss_kernel(sigma=2.0, radius=6):
    x = np.arange(-radius, radius + 1, dtype=np.float64)
    k = np.exp(-(x * x) / (2 * sigma * sigma))
    return k / k.sum()


def _blur(img):
    k = _gauss_kernel()
    r = len(k) // 2
    p = np.pad(img, ((0, 0), (r, r)), mode="reflect")
    out = np.zeros_like(img)
    for i, kv in enumerate(k):
        out += kv * p[:, i:i + img.shape[1]]
    p = np.pad(out, ((r, r), (0, 0)), mode="reflect")
    out2 = np.zeros_like(img)
    for i, kv in enumerate(k):
        out2 += kv * p[i:i + img.shape[0], :]
    return out2


def textured_frame(w, h, seed, contrast=1.0):
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 256, size=(h, w)).astype(np.float64)
    a = _blur(a)
    a = (a - a.min()) / (a.max() - a.min())
    if contrast != 1.0:
        a = 0.5 + (a - 0.5) * contrast
    return np.clip(np.rint(a * 255.0), 0, 255).astype(np.uint8)


def mixed_frame(w, h, seed):
    """Left half full contrast, right half low contrast (fallback cells), a flat band (empty cells)."""
    a = textured_frame(w, h, seed).astype(np.float64)
    b = 128 + (a - 128) * 0.08
    out = a.copy()
    out[:, w // 2:] = b[:, w // 2:]
    out[h // 3: h // 3 + 45, :] = 97
    return np.clip(np.rint(out), 0, 255).astype(np.uint8)


def frame_batch(w, h, n, seed0=0, distinct=8):
    """n frames built from `distinct` seeded textures (rolled so that every frame differs)."""
    base = [textured_frame(w, h, seed0 + i) for i in range(min(distinct, n))]
    out = np.empty((n, h, w), np.uint8)
    for i in range(n):
        out[i] = np.roll(base[i % len(base)], shift=(3 * (i // len(base)), 5 * (i // len(base))), axis=(0, 1))
    return out


def random_descriptors(n, seed):
    rng = np.random.default_rng(seed)
    return rng.integers(0, 256, size=(n, 32), dtype=np.uint8)


def flip_bits(desc, nflips, rng):
    """Return a copy of desc (n,32) with nflips[i] random distinct bit flips in row i."""
    out = desc.copy()
    for i in range(desc.shape[0]):
        k = int(nflips[i])
        if k <= 0:
            continue
        bits = rng.choice(256, size=k, replace=False)
        for b in bits:
            out[i, b >> 3] ^= np.uint8(1 << (b & 7))
    return out


# ---- matcher inputs -------------------------------------------
def scale_tables(scale=1.2, levels=8):
    """mvScaleFactor / mvInvScaleFactor / mvLevelSigma2 as the extractor builds them (float chain)."""
    sf = np.ones(levels, np.float32)
    for i in range(1, levels):
        sf[i] = np.float32(np.float64(sf[i - 1]) * np.float64(np.float32(scale)))
    return sf, (np.float32(1.0) / sf).astype(np.float32), (sf * sf).astype(np.float32)


def projection_case(kps, desc, n_mp, seed, jitter=3.0, p_not_in_view=0.05, p_no_obs=0.03, stereo=False):
    """n_mp synthetic MapPoints drawn from a frame's own keypoints: projected position = keypoint
    +-jitter px, level = source octave, descriptor = source descriptor with 0..40 random bit flips."""
    rng = np.random.default_rng(seed)
    n = len(kps)
    src = rng.integers(0, n, size=n_mp)
    mp = {
        "proj_x": (kps["x"][src] + rng.uniform(-jitter, jitter, n_mp)).astype(np.float32),
        "proj_y": (kps["y"][src] + rng.uniform(-jitter, jitter, n_mp)).astype(np.float32),
        "view_cos": rng.uniform(0.9, 1.0, n_mp).astype(np.float32),
        "level": kps["octave"][src].astype(np.int32),
        "in_view": (rng.uniform(size=n_mp) >= p_not_in_view).astype(np.uint8),
        "has_obs": (rng.uniform(size=n_mp) >= p_no_obs).astype(np.uint8),
        "desc": flip_bits(desc[src], rng.integers(0, 41, size=n_mp), rng),
        "src": src,
    }
    mp["proj_xr"] = (mp["proj_x"] - rng.uniform(2, 60, n_mp)).astype(np.float32)
    u_right = None
    if stereo:
        u_right = np.where(rng.uniform(size=n) < 0.6, kps["x"] - rng.uniform(2, 60, n), -1.0).astype(np.float32)
    claimed = (rng.uniform(size=n) < 0.05).astype(np.uint8)
    return mp, u_right, claimed


def lastframe_case(kps, desc, n_pts, seed, motion=(0.02, -0.01, 0.3), jitter=2.0, stereo=False, fx=458.654, fy=457.296,
                   cx=367.215, cy=248.375, bf=47.9, p_invalid=0.1, p_no_obs=0.05, rot_outliers=0.15):
    """A LastFrame for SearchByProjection(CurrentFrame, LastFrame, ...): n_pts of its keypoints carry MapPoints whose
    world positions project (with the current pose) next to a keypoint of the CURRENT frame (kps/desc), so that the
    windows are populated; some keypoint angles are rotated to feed the rotation histogram with outliers.
    pose = identity current pose shifted by `motion` relative to the last one (tlc = motion)."""
    rng = np.random.default_rng(seed)
    n = len(kps)
    src = rng.integers(0, n, size=n_pts)
    u = kps["x"][src] + rng.uniform(-jitter, jitter, n_pts)
    v = kps["y"][src] + rng.uniform(-jitter, jitter, n_pts)
    z = rng.uniform(1.0, 12.0, n_pts)
    z[rng.uniform(size=n_pts) < 0.02] *= -1.0                     # behind the camera: skipped (:1256)
    th = rng.uniform(-0.02, 0.02, 3)                              # small rotation, exact orthonormality is irrelevant here
    Rcw = np.array([[1, -th[2], th[1]], [th[2], 1, -th[0]], [-th[1], th[0], 1]], np.float64)
    tcw = rng.uniform(-0.2, 0.2, 3)
    pc = np.stack([(u - cx) / fx * z, (v - cy) / fy * z, z], 1)   # camera coordinates
    pw = (np.linalg.inv(Rcw) @ (pc - tcw).T).T                    # world coordinates
    # LastFrame pose: tlc = Rlw * twc + tlw = motion  (with Rlw = Rcw: tlw = motion + tcw)
    Rlw, tlw = Rcw.copy(), np.asarray(motion, np.float64) + tcw
    mb = bf / fx
    pose = np.concatenate([Rcw.reshape(9), tcw, Rlw.reshape(9), tlw, [fx, fy, cx, cy, bf, mb]]).astype(np.float32)
    ang = kps["angle"][src] + rng.normal(0, 3.0, n_pts)
    out = rng.uniform(size=n_pts) < rot_outliers
    ang = np.mod(np.where(out, ang + rng.uniform(40, 320, n_pts), ang), 360.0)
    octave = np.clip(kps["octave"][src] + rng.integers(-1, 2, n_pts), 0, 7)
    last = {
        "valid": (rng.uniform(size=n_pts) >= p_invalid).astype(np.uint8),
        "has_obs": (rng.uniform(size=n_pts) >= p_no_obs).astype(np.uint8),
        "world_pos": pw.astype(np.float32),
        "desc": flip_bits(desc[src], rng.integers(0, 50, size=n_pts), rng),
        "octave": octave.astype(np.int32),
        "angle": ang.astype(np.float32),
        "pose": pose,
        "src": src,
    }
    u_right = None
    if stereo:
        u_right = np.where(rng.uniform(size=n) < 0.6, kps["x"] - bf / rng.uniform(1.0, 12.0, n), -1.0).astype(np.float32)
    claimed = (rng.uniform(size=n) < 0.05).astype(np.uint8)
    return last, u_right, claimed


def keyframe_points_case(kps, desc, n_pts, seed, variant, jitter=2.0, fx=458.654, fy=457.296, cx=367.215, cy=248.375,
                         scale=1.2, nlevels=8, p_invalid=0.1, rot_outliers=0.15):
    """MapPoints for the projection searches with a predicted level (variant: 0 relocalisation, 1 Sim3 SearchByProjection,
    2 Fuse, 3 Fuse with Sim3, 4 one direction of SearchBySim3): world points that project next to keypoints of the
    searched frame (kps/desc), with mfMaxDistance chosen so that PredictScale lands near the source
    keypoint's octave; a few points violate the distance range, the viewing angle (Sim3) or sit behind the camera."""
    rng = np.random.default_rng(seed)
    n = len(kps)
    src = rng.integers(0, n, size=n_pts)
    u = kps["x"][src] + rng.uniform(-jitter, jitter, n_pts)
    v = kps["y"][src] + rng.uniform(-jitter, jitter, n_pts)
    z = rng.uniform(1.0, 12.0, n_pts)
    z[rng.uniform(size=n_pts) < 0.03] *= -1.0
    th = rng.uniform(-0.02, 0.02, 3)
    Rcw = np.array([[1, -th[2], th[1]], [th[2], 1, -th[0]], [-th[1], th[0], 1]], np.float64)
    tcw = rng.uniform(-0.2, 0.2, 3)
    pc = np.stack([(u - cx) / fx * z, (v - cy) / fy * z, z], 1)
    pw = (np.linalg.inv(Rcw) @ (pc - tcw).T).T
    Ow = -Rcw.T @ tcw
    PO = pw - Ow
    dist = np.linalg.norm(PO, axis=1)
    # PredictScale = ceil(log(maxD / dist) / log(scale)): aim at the source octave (+-1), fractional part uniform
    lvl = np.clip(kps["octave"][src] + rng.integers(-1, 2, n_pts), 0, nlevels - 1)
    max_d = dist * scale ** (lvl - rng.uniform(0.05, 0.95, n_pts))
    min_d = max_d / scale ** (nlevels - 1)
    bad_range = rng.uniform(size=n_pts) < 0.05
    max_d = np.where(bad_range, dist * 0.5, max_d)
    normal = PO / dist[:, None] + rng.normal(0, 0.2, (n_pts, 3))
    steep = rng.uniform(size=n_pts) < 0.08
    normal[steep] *= -1.0
    ang = kps["angle"][src] + rng.normal(0, 3.0, n_pts)
    out = rng.uniform(size=n_pts) < rot_outliers
    ang = np.mod(np.where(out, ang + rng.uniform(40, 320, n_pts), ang), 360.0)
    variant = int(variant)
    pose = np.zeros(30, np.float32)
    pose[:9] = Rcw.reshape(9); pose[9:12] = tcw; pose[24:28] = [fx, fy, cx, cy]
    bf = 47.9
    if variant in (1, 2, 3):
        pose[12:15] = Ow
        pose[28] = bf
    u_right = np.where(rng.uniform(size=n) < 0.5, kps["x"] - bf / rng.uniform(1.0, 12.0, n), -1.0).astype(np.float32)
    if variant == 2:   # stereo keypoints consistent with their source points, so that the chi-square gate passes some
        zs = np.full(n, np.nan); zs[src] = z
        u_right = np.where(np.isfinite(zs) & (rng.uniform(size=n) < 0.7), kps["x"] - bf / np.abs(zs) + rng.normal(0, 0.7, n),
                           u_right).astype(np.float32)
    if variant == 4:
        # world -> camera 1 (R1w, t1w) -> camera 2 (sR21, t21), composed to the Rcw, tcw used above; distances are measured
        # in camera 2, so that is where max_d has to refer to
        a = rng.uniform(-0.3, 0.3, 3)
        R1w = np.array([[1, -a[2], a[1]], [a[2], 1, -a[0]], [-a[1], a[0], 1]], np.float64)
        t1w = rng.uniform(-1, 1, 3)
        sR21 = Rcw @ np.linalg.inv(R1w)
        t21 = tcw - sR21 @ t1w
        pose[:9] = R1w.reshape(9); pose[9:12] = t1w; pose[12:21] = sR21.reshape(9); pose[21:24] = t21
        dist2 = np.linalg.norm(pc, axis=1)
        max_d = np.where(bad_range, dist2 * 0.5, dist2 * scale ** (lvl - rng.uniform(0.05, 0.95, n_pts)))
        min_d = max_d / scale ** (nlevels - 1)
    pts = {
        "valid": (rng.uniform(size=n_pts) >= p_invalid).astype(np.uint8),
        "world_pos": pw.astype(np.float32), "normal": normal.astype(np.float32),
        "desc": flip_bits(desc[src], rng.integers(0, 45, size=n_pts), rng),
        "angle": ang.astype(np.float32), "max_distance": max_d.astype(np.float32), "min_distance": min_d.astype(np.float32),
        "pose": pose, "log_scale_factor": float(np.float32(np.log(np.float32(scale)))), "n_scale_levels": nlevels, "src": src,
        "u_right": u_right,
    }
    claimed = (rng.uniform(size=n) < 0.05).astype(np.uint8)
    return pts, claimed


def sim3_pair_case(kps, desc, seed, s12=1.3, jitter=1.0, fx=458.654, fy=457.296, cx=367.215, cy=248.375, scale=1.2, nlevels=8,
                   p_invalid=0.1):
    """A consistent input of ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th) (src/ORBmatcher.cc:1006-1216):
    two keyframes that share the given keypoints, one MapPoint per keypoint on either side, MapPoint i of KF1 projecting
    (through R1w, t1w and then sR21, t21) next to keypoint i of KF2 and vice versa, so that mutual matches are common.
    Returns dict: pose1 / pose2 (Rkw 9 + tkw 3), s12, R12, t12, pts1, pts2 (valid, world_pos, desc, max_distance, min_distance,
    normal, angle), intr, already12."""
    rng = np.random.default_rng(seed)
    n = len(kps)

    def small_rot(a):
        return np.array([[1, -a[2], a[1]], [a[2], 1, -a[0]], [-a[1], a[0], 1]], np.float64)
    R1w, t1w = small_rot(rng.uniform(-0.3, 0.3, 3)), rng.uniform(-1, 1, 3)
    R2w, t2w = small_rot(rng.uniform(-0.3, 0.3, 3)), rng.uniform(-1, 1, 3)
    R12 = small_rot(rng.uniform(-0.05, 0.05, 3))
    t12 = rng.uniform(-0.3, 0.3, 3)
    sR12 = s12 * R12
    sR21 = (1.0 / s12) * R12.T
    t21 = -sR21 @ t12

    def side(Rw, tw, A, b):
        # target camera coordinates next to keypoint i of the OTHER keyframe: pc_other = A (Rw pw + tw) + b
        u = kps["x"] + rng.uniform(-jitter, jitter, n)
        v = kps["y"] + rng.uniform(-jitter, jitter, n)
        z = rng.uniform(1.0, 12.0, n)
        z[rng.uniform(size=n) < 0.03] *= -1.0
        pc = np.stack([(u - cx) / fx * z, (v - cy) / fy * z, z], 1)
        pw = (np.linalg.inv(Rw) @ ((np.linalg.inv(A) @ (pc - b).T).T - tw).T).T
        dist = np.linalg.norm(pc, axis=1)
        lvl = np.clip(kps["octave"] + rng.integers(-1, 2, n), 0, nlevels - 1)
        max_d = dist * scale ** (lvl - rng.uniform(0.05, 0.95, n))
        max_d = np.where(rng.uniform(size=n) < 0.05, dist * 0.5, max_d)
        return {"valid": (rng.uniform(size=n) >= p_invalid).astype(np.uint8), "world_pos": pw.astype(np.float32),
                "desc": flip_bits(desc, rng.integers(0, 45, size=n), rng), "max_distance": max_d.astype(np.float32),
                "min_distance": (max_d / scale ** (nlevels - 1)).astype(np.float32), "normal": np.zeros((n, 3), np.float32),
                "angle": kps["angle"].astype(np.float32), "log_scale_factor": float(np.float32(np.log(np.float32(scale)))),
                "n_scale_levels": nlevels}
    pts1 = side(R1w, t1w, sR21, t21)        # KF1's points are searched in KF2
    pts2 = side(R2w, t2w, sR12, t12)
    already = np.where(rng.uniform(size=n) < 0.05, rng.integers(0, n, n), -1).astype(np.int32)
    f = lambda a: np.asarray(a, np.float32).reshape(-1)
    return {"pose1": np.concatenate([f(R1w), f(t1w)]), "pose2": np.concatenate([f(R2w), f(t2w)]), "s12": float(np.float32(s12)),
            "R12": f(R12), "t12": f(t12), "pts1": pts1, "pts2": pts2, "intr": np.array([fx, fy, cx, cy], np.float32),
            "already12": already}


def initialization_case(kps, desc, seed, shift=(12.0, -7.0), n_dup=200):
    """Two frames for SearchForInitialization: F1 = the given features (60 % forced to octave 0) plus n_dup near-copies
    that compete for the same partner; F2 = shifted, jittered, bit-flipped, shuffled counterparts of 80 % of F1 with some
    rotated angles.  Returns (kps1, desc1, kps2, desc2, prev_matched)."""
    rng = np.random.default_rng(seed)
    k1 = kps.copy()
    k1["octave"] = np.where(rng.uniform(size=len(k1)) < 0.6, 0, k1["octave"])
    dup = rng.integers(0, len(k1), n_dup)
    kd = k1[dup].copy()
    kd["x"] += rng.uniform(-3, 3, n_dup).astype(np.float32); kd["y"] += rng.uniform(-3, 3, n_dup).astype(np.float32)
    kd["octave"] = 0
    k1 = np.concatenate([k1, kd]); d1 = np.concatenate([desc, flip_bits(desc[dup], rng.integers(0, 12, n_dup), rng)])
    order = rng.permutation(len(k1))
    k1, d1 = k1[order], d1[order]
    sel = np.flatnonzero(rng.uniform(size=len(kps)) < 0.8)
    k2 = kps[sel].copy()
    k2["octave"] = np.where(rng.uniform(size=len(k2)) < 0.75, 0, k2["octave"])
    k2["x"] = (k2["x"] + shift[0] + rng.uniform(-2, 2, len(k2))).astype(np.float32)
    k2["y"] = (k2["y"] + shift[1] + rng.uniform(-2, 2, len(k2))).astype(np.float32)
    ang = k2["angle"] + rng.normal(0, 3.0, len(k2))
    ang = np.where(rng.uniform(size=len(k2)) < 0.12, ang + rng.uniform(40, 320, len(k2)), ang)
    k2["angle"] = np.mod(ang, 360.0).astype(np.float32)
    d2 = flip_bits(desc[sel], rng.integers(0, 40, len(sel)), rng)
    o2 = rng.permutation(len(k2))
    k2, d2 = k2[o2], d2[o2]
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    return k1, d1, k2, d2, prev


def feature_vector(n, n_nodes, seed, base=None, p_keep=0.8):
    """Synthetic DBoW2::FeatureVector as CSR: ascending node ids, features in insertion (index) order.
    With `base` (node of each feature of another keyframe mapped through `src`) matched features
    share nodes with probability p_keep."""
    rng = np.random.default_rng(seed)
    node = rng.integers(0, n_nodes, size=n) * 7 + 3
    if base is not None:
        keep = rng.uniform(size=n) < p_keep
        node = np.where(keep, base, node)
    ids = np.unique(node)
    start = np.zeros(len(ids) + 1, np.int32)
    feats = []
    for k, nid in enumerate(ids):
        f = np.nonzero(node == nid)[0]
        feats.append(f)
        start[k + 1] = start[k] + len(f)
    feats = np.concatenate(feats).astype(np.uint32) if feats else np.zeros(0, np.uint32)
    return node, ids.astype(np.uint32), start, feats


def triangulation_case(kps, desc, seed, levels=8, scale=1.2, n_nodes=120):
    """KF1 = the given features; KF2 = a shuffled subset moved along x (pure x translation, so the
    epipolar line of (x1, y1) is y2 ~= y1) with descriptor noise, plus unrelated features."""
    rng = np.random.default_rng(seed)
    n1 = len(kps)
    sf, isf, s2 = scale_tables(scale, levels)
    perm = rng.permutation(n1)
    n2 = n1
    kps2 = kps[perm].copy()
    moved = rng.uniform(size=n2) < 0.7
    kps2["x"] = np.where(moved, kps2["x"] - rng.uniform(1, 40, n2), rng.uniform(20, 700, n2)).astype(np.float32)
    kps2["y"] = np.where(moved, kps2["y"] + rng.normal(0, 1.0, n2) * sf[kps2["octave"]], rng.uniform(20, 450, n2)).astype(np.float32)
    kps2["angle"] = np.mod(kps2["angle"] + rng.normal(0, 8, n2), 360).astype(np.float32)
    desc2 = flip_bits(desc[perm], np.where(moved, rng.integers(0, 60, n2), 128), rng)
    node1, ids1, st1, f1 = feature_vector(n1, n_nodes, seed + 1)
    node2, ids2, st2, f2 = feature_vector(n2, n_nodes, seed + 2, base=node1[perm])
    kf1 = {"kps": kps, "desc": desc, "u_right": np.where(rng.uniform(size=n1) < 0.3, kps["x"] - 5, -1).astype(np.float32),
           "has_mp": (rng.uniform(size=n1) < 0.3).astype(np.uint8), "node_ids": ids1, "node_start": st1, "node_feats": f1,
           "scale_factors": sf, "level_sigma2": s2}
    kf2 = {"kps": kps2, "desc": desc2, "u_right": np.where(rng.uniform(size=n2) < 0.3, kps2["x"] - 5, -1).astype(np.float32),
           "has_mp": (rng.uniform(size=n2) < 0.3).astype(np.uint8), "node_ids": ids2, "node_start": st2, "node_feats": f2,
           "scale_factors": sf, "level_sigma2": s2}
    F12 = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32) + rng.normal(0, 1e-5, (3, 3)).astype(np.float32)
    return kf1, kf2, F12, np.float32(400.0), np.float32(240.0)


def stereo_pair(w, h, seed, max_disp=40):
    """Left = textured frame; right = left resampled with a smooth disparity field (2..max_disp px) + noise."""
    rng = np.random.default_rng(seed)
    wide = textured_frame(w + max_disp + 8, h, seed).astype(np.float64)
    left = wide[:, :w]
    disp = 2 + (max_disp - 2) * (0.5 + 0.5 * np.sin(np.linspace(0, 3.0, h)))[:, None] * np.ones((1, w))
    xs = np.arange(w)[None, :] + disp       # right(x) = left(x + d): a point at uL appears at uR = uL - d
    x0 = np.floor(xs).astype(int)
    fr = xs - x0
    rows = np.arange(h)[:, None]
    right = wide[rows, x0] * (1 - fr) + wide[rows, x0 + 1] * fr + rng.normal(0, 2.0, (h, w))
    return np.clip(np.rint(left), 0, 255).astype(np.uint8), np.clip(np.rint(right), 0, 255).astype(np.uint8)


def knn_case(nq, nt, seed):
    """Config 5: train uniform random; half of the queries are train rows with Binomial(256, 0.1) bit flips."""
    rng = np.random.default_rng(seed)
    train = rng.integers(0, 256, size=(nt, 32), dtype=np.uint8)
    query = rng.integers(0, 256, size=(nq, 32), dtype=np.uint8)
    half = nq // 2
    src = rng.integers(0, nt, size=half)
    noise = np.packbits((rng.uniform(size=(half, 256)) < 0.1).astype(np.uint8), axis=1, bitorder="little")
    query[:half] = train[src] ^ noise
    return query, train


def undistort_maps(w, h, k1=-0.28, k2=0.07):
    """Fixed-point remap maps of the CV_16SC2 / CV_16UC1 kind (integer source position + 5-bit fractions packed as
    fy * 32 + fx) for a radial distortion model; part of the border maps outside the source image."""
    fx, fy, cx, cy = 0.61 * w, 0.95 * h, 0.49 * w, 0.52 * h
    xs, ys = np.meshgrid(np.arange(w, dtype=np.float64), np.arange(h, dtype=np.float64))
    x, y = (xs - cx) / fx, (ys - cy) / fy
    r2 = x * x + y * y
    d = 1 + k1 * r2 + k2 * r2 * r2
    u, v = x * d * fx + cx, y * d * fy + cy
    iu, iv = np.round(u * 32).astype(np.int64), np.round(v * 32).astype(np.int64)
    map1 = np.stack([iu >> 5, iv >> 5], -1).astype(np.int16)
    map2 = ((iv & 31) * 32 + (iu & 31)).astype(np.uint16)
    return map1, map2


def quat_from_rotvec(r):
    """Unit quaternion (x, y, z, w) of a rotation vector, float32."""
    r = np.asarray(r, np.float64)
    th = np.linalg.norm(r)
    if th < 1e-12:
        return np.array([0, 0, 0, 1], np.float32)
    ax = r / th
    return np.concatenate([ax * np.sin(th / 2), [np.cos(th / 2)]]).astype(np.float32)


def quat_rotate(q, v):
    """R(q) v in float64 (plain rotation-matrix form; an independent route from Eigen's _transformVector)."""
    x, y, z, w = (float(a) for a in q)
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                  [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                  [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
    return R @ np.asarray(v, np.float64)


def pyramid_of(img, scale=1.2, levels=8):
    """mvImagePyramid as ORBextractor::ComputePyramid builds it (sizes from level 0, each level resized from the previous)."""
    import cv2
    out = [np.ascontiguousarray(img)]
    sc, inv, _ = scale_tables(scale, levels)
    for l in range(1, levels):
        w = int(np.rint(np.float32(img.shape[1]) * inv[l])); h = int(np.rint(np.float32(img.shape[0]) * inv[l]))
        out.append(cv2.resize(out[-1], (w, h), interpolation=cv2.INTER_LINEAR))
    return out


def direct_projection_case(seed, n=300, w=752, h=480, fx=458.654, fy=457.296, cx=367.215, cy=248.375):
    """Two reference keyframes and a current frame for FindDirectProjection: the current image is keyframe 0's image shifted by
    the parallax of a sideways camera motion at the common depth of most points, so Align2D converges for them; other points sit
    at wrong depths, near the border, or belong to keyframe 1 (different intrinsics: the affine warp picks search level > 0)."""
    import cv2
    rng = np.random.default_rng(seed)
    ref = textured_frame(w, h, 40 + seed)
    Z, t_cr = 5.0, np.array([0.06, -0.035, 0.0])
    shift = np.array([fx * t_cr[0] / Z, fy * t_cr[1] / Z])
    M = np.array([[1, 0, shift[0]], [0, 1, shift[1]]], np.float64)
    cur = cv2.warpAffine(ref, M, (w, h), flags=cv2.INTER_LINEAR, borderMode=cv2.BORDER_REFLECT_101)
    ref2 = textured_frame(w, h, 90 + seed)
    sc, inv, _ = scale_tables()
    q_ref = quat_from_rotvec([0.05, -0.08, 0.02]); t_ref = np.array([0.3, -0.2, 0.1], np.float32)
    q_cr = quat_from_rotvec([0.0, 0.0, 0.004]); t_cr32 = t_cr.astype(np.float32)
    kfs = [dict(levels=pyramid_of(ref), pose_ref=np.concatenate([q_ref, t_ref]), t_cur_ref=np.concatenate([q_cr, t_cr32]),
                fx=fx, fy=fy, cx=cx, cy=cy),
           dict(levels=pyramid_of(ref2), pose_ref=np.concatenate([q_ref, t_ref]), t_cur_ref=np.concatenate([q_cr, t_cr32]),
                fx=fx / 2.2, fy=fy / 2.2, cx=cx, cy=cy)]
    q_inv = q_ref * np.array([-1, -1, -1, 1], np.float32)
    kf_index = (rng.random(n) < 0.15).astype(np.int32)
    octave = rng.choice(8, size=n, p=[.3, .2, .15, .1, .1, .05, .05, .05]).astype(np.int32)
    px_ref = np.stack([rng.uniform(30, w - 30, n), rng.uniform(30, h - 30, n)], axis=1).astype(np.float32)
    px_ref[: n // 20] = np.stack([rng.uniform(0, 12, n // 20), rng.uniform(0, h, n // 20)], axis=1)   # warp leaves the image
    depth = np.where(rng.random(n) < 0.8, Z, rng.uniform(1.0, 20.0, n))
    world = np.zeros((n, 3), np.float32); px_curr = np.zeros((n, 2), np.float32)
    for i in range(n):
        k = kfs[kf_index[i]]
        p_ref = np.array([(px_ref[i, 0] - k["cx"]) * depth[i] / k["fx"], (px_ref[i, 1] - k["cy"]) * depth[i] / k["fy"], depth[i]])
        world[i] = quat_rotate(q_inv, p_ref - t_ref.astype(np.float64))
        p_cur = quat_rotate(q_cr, p_ref) + t_cr
        px_curr[i] = [fx * p_cur[0] / p_cur[2] + cx, fy * p_cur[1] / p_cur[2] + cy]
    px_curr += rng.uniform(-1.5, 1.5, size=(n, 2)).astype(np.float32)
    px_curr[n // 20: n // 10] = np.stack([rng.uniform(w - 6, w + 5, n // 10 - n // 20), rng.uniform(0, h, n // 10 - n // 20)], axis=1)
    return dict(kfs=kfs, cur_levels=pyramid_of(cur), cur_intr=np.array([fx, fy, cx, cy], np.float32), scale=sc, inv_scale=inv,
                inv_sigma2_1=float(np.float32(1.0) / (sc[1] * sc[1])), kf_index=kf_index, px_ref=px_ref, octave=octave,
                world_pos=world, px_curr=px_curr)


def sparse_align_case(seed, n=800, w=752, h=480, fx=458.654, fy=457.296, cx=367.215, cy=248.375, perturb=(0.02, -0.012, 0.0)):
    """Reference + current frame for SparseImgAlign: the current image is the reference shifted by the parallax of a sideways
    motion t_true at the common depth Z of the features (a fronto-parallel plane), so the optimum of the alignment is known; the
    initial guess is off by `perturb`.  Some features are invalid (no MapPoint / outlier), some too close to the border."""
    import cv2
    rng = np.random.default_rng(seed)
    ref = textured_frame(w, h, 60 + seed)
    Z, t_true = 4.0, np.array([0.05, -0.03, 0.0])
    shift = np.array([fx * t_true[0] / Z, fy * t_true[1] / Z])
    M = np.array([[1, 0, shift[0]], [0, 1, shift[1]]], np.float64)
    cur = cv2.warpAffine(ref, M, (w, h), flags=cv2.INTER_LINEAR, borderMode=cv2.BORDER_REFLECT_101)
    sc, inv, _ = scale_tables()
    q_ref = quat_from_rotvec([0.03, 0.05, -0.02]); t_ref = np.array([-0.2, 0.1, 0.3], np.float32)
    q_inv = q_ref * np.array([-1, -1, -1, 1], np.float32)
    kp = np.stack([rng.uniform(2, w - 2, n), rng.uniform(2, h - 2, n)], axis=1).astype(np.float32)
    valid = (rng.random(n) < 0.9).astype(np.uint8)
    world = np.zeros((n, 3), np.float32)
    for i in range(n):
        p_ref = np.array([(kp[i, 0] - cx) * Z / fx, (kp[i, 1] - cy) * Z / fy, Z])
        world[i] = quat_rotate(q_inv, p_ref - t_ref.astype(np.float64))
    t0 = np.concatenate([quat_from_rotvec([0.001, -0.0015, 0.002]), (t_true + np.array(perturb)).astype(np.float32)])
    return dict(ref_levels=pyramid_of(ref), cur_levels=pyramid_of(cur), inv_scale=inv, focal=fx,
                cur_intr=np.array([fx, fy, cx, cy], np.float32), ref_pose=np.concatenate([q_ref, t_ref]), kp_xy=kp, valid=valid,
                world_pos=world, t_cur_ref=t0.astype(np.float32), t_true=t_true)
