"""The matcher parity cases, shared by
  * tools/make_golden_match.py  — runs the REFERENCE'S OWN code (oracle/_ref/libygz_ref_match.so) on them and commits its
                                  outputs as tests/golden/match_golden.npz,
  * tests/test_ref_match.py     — CPU: restatement (oracle/match_oracle.cpp, direct_oracle.cpp) == golden == live reference,
  * tests/test_gpu_match_golden.py — GPU: kernels through the C ABI == golden.
A case is (kind, key, params); build_inputs() regenerates its inputs from seeds (tests/synth.py); inputs that only the
reference can derive (the decomposition of a Sim3 matrix, TCR) are stored in the golden file next to the outputs.
Every run_* returns a list of numpy arrays in a fixed order; floats are compared by their bytes."""
import functools

import numpy as np

import oracle_lib as O
import synth

BOUNDS = (0.0, 0.0, 752.0, 480.0)


@functools.lru_cache(maxsize=None)
def feats():
    img = synth.textured_frame(752, 480, 0)
    kps, desc = O.extract(O.params(1000), img)
    return img, kps, desc


@functools.lru_cache(maxsize=None)
def tables():
    sf, isf, s2 = synth.scale_tables()
    return sf, isf, s2, (np.float32(1.0) / s2).astype(np.float32)


def _grid(*lists):
    return [tuple(x) for x in lists]


# parameter lists = those of tests/test_gpu_match.py / test_gpu_direct.py (BASELINE configs 2 and 3 shapes) + edge cases
CASES = []
for th, cl, st, seed in [(1.0, False, False, 1), (5.0, False, True, 2), (3.0, True, False, 3), (15.0, False, False, 4), (8.0, True, True, 5)]:
    for k in range(3):
        CASES.append(("sbp", "sbp/th%g_cl%d_st%d_s%d_k%d" % (th, cl, st, seed, k), (th, cl, st, seed, k)))
CASES.append(("sbp_chain", "sbp/claim_chain", ()))
CASES.append(("sbp_empty", "sbp/empty", ()))
for th, mono, cl, co, st, motion, seed in [
        (15.0, True, False, True, False, (0.02, -0.01, 0.3), 1), (7.0, False, True, True, True, (0.0, 0.0, 0.5), 2),
        (7.0, False, True, False, True, (0.0, 0.0, -0.5), 3), (15.0, False, True, True, False, (0.05, 0.0, 0.01), 4),
        (40.0, True, False, True, False, (0.0, 0.0, 0.0), 5)]:
    for k in range(3):
        CASES.append(("last", "last/th%g_m%d_cl%d_co%d_st%d_s%d_k%d" % (th, mono, cl, co, st, seed, k), (th, mono, cl, co, st, motion, seed, k)))
for th, variant, od, co, seed in [(10.0, 0, 100, True, 1), (3.0, 0, 64, True, 2), (15.0, 0, 100, False, 3), (10.0, 1, 50, False, 4),
                                  (4.0, 1, 50, False, 5), (40.0, 1, 50, False, 6), (3.0, 2, 50, False, 7), (2.5, 2, 50, False, 8),
                                  (4.0, 3, 50, False, 9), (20.0, 3, 50, False, 10)]:
    for k in range(3):
        CASES.append(("kf", "kf/v%d_th%g_od%d_co%d_s%d_k%d" % (variant, th, od, co, seed, k), (th, variant, od, co, seed, k)))
for th, seed, s12 in [(7.5, 1, 1.3), (3.0, 2, 0.8), (10.0, 3, 1.0)]:
    CASES.append(("sim3", "sim3/th%g_s%d" % (th, seed), (th, seed, s12)))
for window, ratio, co, seed in [(100, 0.9, True, 1), (30, 0.9, True, 2), (100, 0.7, False, 3), (400, 0.95, True, 4)]:
    for k in range(3):
        CASES.append(("init", "init/w%d_r%g_co%d_s%d_k%d" % (window, ratio, co, seed, k), (window, ratio, co, seed, k)))
for only_stereo, co, seed in [(False, False, 1), (False, True, 2), (True, True, 3)]:
    CASES.append(("tri", "tri/os%d_co%d_s%d" % (only_stereo, co, seed), (only_stereo, co, seed)))
for frame2, co, ratio, nn, seed in [(True, True, 0.7, 100, 1), (True, False, 0.9, 30, 2), (False, True, 0.75, 100, 3), (False, False, 0.6, 5, 4)]:
    CASES.append(("bow", "bow/f%d_co%d_r%g_n%d_s%d" % (frame2, co, ratio, nn, seed), (frame2, co, ratio, nn, seed)))
for w, h, nf, seed in [(640, 240, 800, 5), (1241, 376, 2000, 6)]:
    CASES.append(("stereo", "stereo/%dx%d_n%d_s%d" % (w, h, nf, seed), (w, h, nf, seed)))
for n_points, seed, big in [(1, 0, None), (7, 1, None), (500, 2, None), (40, 3, 700)]:
    CASES.append(("distinct", "distinct/n%d_s%d" % (n_points, seed), (n_points, seed, big)))
for seed, n in [(0, 1), (1, 33), (2, 400), (3, 3000)]:
    CASES.append(("fdp", "fdp/s%d_n%d" % (seed, n), (seed, n)))
# SparseImgAlign (src/SparseImageAlign.cc + include/NLSSolver_impl.hpp): one precompute + computeResiduals at a level ("sares"),
# and the whole coarse-to-fine run ("sarun").  Levels <= 5 only: run() reads `int iterations[] = {10, 10, 10, 10, 10, 10}` at
# index level_ (:37, :42), so with 8 pyramid levels the reference reads past the array at levels 6 and 7 (undefined behaviour;
# the compiled reference here happens to find 0 there and skips both levels).  DESIGN.md 2 records it.
for seed, n, level in [(0, 300, 1), (1, 300, 3), (2, 800, 5), (3, 50, 2)]:
    CASES.append(("sares", "sares/s%d_n%d_l%d" % (seed, n, level), (seed, n, level)))
for seed, n, hi, lo, all_invalid in [(0, 300, 5, 1, False), (1, 800, 5, 1, False), (2, 300, 4, 2, False), (3, 40, 5, 1, False), (4, 8, 5, 1, True)]:
    CASES.append(("sarun", "sarun/s%d_n%d_l%d_%d_inv%d" % (seed, n, hi, lo, all_invalid), (seed, n, hi, lo, all_invalid)))
GPU_TOLERANCE_KINDS = ("sares", "sarun")   # sums over features: sequential float in the reference, a tree on the device

BY_KEY = {c[1]: c for c in CASES}


def _qmul(a, b):   # quaternions x, y, z, w
    ax, ay, az, aw = [float(v) for v in a]
    bx, by, bz, bw = [float(v) for v in b]
    return np.array([aw * bx + ax * bw + ay * bz - az * by, aw * by + ay * bw + az * bx - ax * bz,
                     aw * bz + az * bw + ax * by - ay * bx, aw * bw - ax * bx - ay * by - az * bz])


def build_inputs(case):
    """Seeded inputs of a case (dict).  Keys starting with 'ref_' are filled by the reference (see derive_with_ref)."""
    kind, key, p = case
    img, kps, desc = feats()
    sf, isf, s2, inv_s2 = tables()
    d = {"kind": kind}
    if kind == "sbp":
        th, cl, st, seed, k = p
        mp, ur, claimed = synth.projection_case(kps, desc, 3000 - 500 * k, seed * 10 + k, stereo=st)
        d.update(kps=kps, desc=desc, u_right=ur, claimed=claimed, mp=mp, th=th, check_level=cl, ratio=0.8)
    elif kind == "sbp_chain":
        rng = np.random.default_rng(5)
        src = np.repeat(rng.integers(0, len(kps), size=40), 25)
        mp = {"proj_x": kps["x"][src].copy(), "proj_y": kps["y"][src].copy(), "proj_xr": np.zeros(len(src), np.float32),
              "view_cos": np.full(len(src), 0.95, np.float32), "level": kps["octave"][src].astype(np.int32),
              "in_view": np.ones(len(src), np.uint8), "has_obs": (rng.uniform(size=len(src)) < 0.8).astype(np.uint8),
              "desc": synth.flip_bits(desc[src], rng.integers(0, 10, size=len(src)), rng)}
        d.update(kind="sbp", kps=kps, desc=desc, u_right=None, claimed=None, mp=mp, th=8.0, check_level=False, ratio=0.9)
    elif kind == "sbp_empty":
        mp, _, _ = synth.projection_case(kps, desc, 50, 1)
        mp = {k: v[:0] for k, v in mp.items()}
        d.update(kind="sbp", kps=kps, desc=desc, u_right=None, claimed=None, mp=mp, th=1.0, check_level=False, ratio=0.8)
    elif kind == "last":
        th, mono, cl, co, st, motion, seed, k = p
        last, ur, claimed = synth.lastframe_case(kps, desc, 1000 - 150 * k, seed * 10 + k, motion=motion, stereo=st)
        d.update(kps=kps, desc=desc, u_right=ur, claimed=claimed, last=last, th=th, mono=mono, check_level=cl, check_ori=co)
    elif kind == "kf":
        th, variant, od, co, seed, k = p
        pts, claimed = synth.keyframe_points_case(kps, desc, 1200 - 200 * k, seed * 10 + k, variant)
        d.update(kps=kps, desc=desc, claimed=claimed, pts=pts, th=th, variant=variant, orb_dist=od, check_ori=co,
                 u_right=pts["u_right"] if variant == 2 else None, inv_sigma2=inv_s2, scw_scale=1.0 + 0.1 * k)
    elif kind == "sim3":
        th, seed, s12 = p
        d.update(kps=kps, desc=desc, th=th, c=synth.sim3_pair_case(kps, desc, seed, s12=s12))
    elif kind == "init":
        window, ratio, co, seed, k = p
        n = len(kps) - 100 * k
        k1, d1, k2, d2, prev = synth.initialization_case(kps[:n], desc[:n], seed * 10 + k)
        d.update(k1=k1, d1=d1, k2=k2, d2=d2, prev=prev, window=window, ratio=ratio, check_ori=co)
    elif kind == "tri":
        only_stereo, co, seed = p
        kf1, kf2, F12, ex, ey = synth.triangulation_case(kps, desc, seed)
        d.update(kf1=kf1, kf2=kf2, F12=F12, ex=ex, ey=ey, only_stereo=only_stereo, check_ori=co)
    elif kind == "bow":
        frame2, co, ratio, nn, seed = p
        kf1, kf2, _, _, _ = synth.triangulation_case(kps, desc, seed, n_nodes=nn)
        kf1["has_mp"] = 1 - kf1["has_mp"]; kf2["has_mp"] = 1 - kf2["has_mp"]
        d.update(kf1=kf1, kf2=kf2, frame2=frame2, ratio=ratio, check_ori=co)
    elif kind == "stereo":
        w, h, nf, seed = p
        left, right = synth.stereo_pair(w, h, seed)
        prm = O.params(nf)
        kl, dl = O.extract(prm, left)
        kr, dr = O.extract(prm, right)
        d.update(left=left, right=right, nf=nf, kl=kl, dl=dl, kr=kr, dr=dr, pyr_l=O.pyramid(prm, left), pyr_r=O.pyramid(prm, right),
                 bf=386.1448, b=386.1448 / 718.856)
    elif kind == "distinct":
        from test_direct_oracle import synth_observations
        n_points, seed, big = p
        od, start = synth_observations(n_points, seed, big=big)
        d.update(obs_desc=od, obs_start=start)
    elif kind == "fdp":
        seed, n = p
        c = synth.direct_projection_case(seed, n=n)
        k0 = c["kfs"][0]
        q = _qmul(k0["t_cur_ref"][:4], k0["pose_ref"][:4])
        q /= np.linalg.norm(q)
        t = synth.quat_rotate(k0["t_cur_ref"][:4].astype(np.float64), k0["pose_ref"][4:].astype(np.float64)) + k0["t_cur_ref"][4:]
        d.update(c=c, cur_pose=np.concatenate([q, t]).astype(np.float32), inv_sigma2=inv_s2)
    elif kind in ("sares", "sarun"):
        seed, n = p[0], p[1]
        c = synth.sparse_align_case(seed, n=n)
        if kind == "sarun" and p[4]:
            c["valid"][:] = 0
        q = _qmul(c["t_cur_ref"][:4], c["ref_pose"][:4])
        q /= np.linalg.norm(q)
        t = synth.quat_rotate(c["t_cur_ref"][:4].astype(np.float64), c["ref_pose"][4:].astype(np.float64)) + c["t_cur_ref"][4:]
        d.update(c=c, cur_pose=np.concatenate([q, t]).astype(np.float32))
        if kind == "sares":
            d.update(level=p[2])
        else:
            d.update(max_level=p[2], min_level=p[3])
    else:
        raise KeyError(kind)
    return d


# ---- the reference's own code (needs oracle/_ref/libygz_ref_match.so) ------------------------------------------------------
def run_ref(d):
    """Returns (outputs, derived): derived = inputs only the reference can produce (dict of arrays, may be empty)."""
    import ref_match_lib as R
    sf, isf, s2, inv_s2 = tables()
    k = d["kind"]
    if k == "sbp":
        nm, a = R.search_by_projection(d["kps"], d["desc"], d["u_right"], d["claimed"], BOUNDS, sf, d["mp"], d["th"], d["check_level"], d["ratio"])
        return [np.int32(nm), a], {}
    if k == "last":
        nm, a = R.search_by_projection_last(d["kps"], d["desc"], d["u_right"], d["claimed"], BOUNDS, sf, d["last"], d["th"], d["mono"],
                                            d["check_level"], d["check_ori"])
        return [np.int32(nm), a], {}
    if k == "kf":
        v, pts, der, scw = d["variant"], d["pts"], {}, None
        if v in (1, 3):
            scw = R.scw_of_pose(pts["pose"], d["scw_scale"])
            der["pose15"] = R.decompose_scw(scw)
            pts = dict(pts, pose=np.concatenate([der["pose15"], pts["pose"][15:]]).astype(np.float32))
        nm, a, b = R.search_by_projection_keyframe(d["kps"], d["desc"], d["claimed"], BOUNDS, sf, pts, d["th"], v, d["orb_dist"],
                                                   d["check_ori"], d["u_right"], d["inv_sigma2"], scw)
        return [np.int32(nm), a] + ([b] if b is not None else []), der
    if k == "sim3":
        c = d["c"]
        nf, m12, der = R.search_by_sim3(d["kps"], d["desc"], d["kps"], d["desc"], BOUNDS, sf, c["intr"], c["pose1"], c["pose2"], c["pts1"],
                                        c["pts2"], c["s12"], c["R12"], c["t12"], d["th"], c["already12"])
        return [np.int32(nf), m12], {"derived": der}
    if k == "init":
        nm, m12, prev = R.search_for_initialization(d["k1"], d["d1"], d["k2"], d["d2"], BOUNDS, d["prev"], d["window"], d["ratio"], d["check_ori"])
        return [np.int32(nm), m12, prev], {}
    if k == "tri":
        nm, m = R.search_for_triangulation(d["kf1"], d["kf2"], d["F12"], d["ex"], d["ey"], d["only_stereo"], d["check_ori"])
        return [np.int32(nm), m], {}
    if k == "bow":
        nm, m = R.search_by_bow(d["kf1"], d["kf2"], d["frame2"], d["ratio"], d["check_ori"])
        return [np.int32(nm), m], {}
    if k == "stereo":
        ur, dp = R.stereo_match(d["kl"], d["dl"], d["kr"], d["dr"], d["pyr_l"], d["pyr_r"], sf, isf, d["bf"], d["b"])
        return [ur, dp], {}
    if k == "distinct":
        ch, out = R.distinctive_descriptors(d["obs_desc"], d["obs_start"])
        return [ch, out * ch[:, None]], {}
    if k == "fdp":
        c = d["c"]
        px, lv, ok, tcr = R.find_direct_projection(c["kfs"], c["cur_levels"], c["cur_intr"], d["cur_pose"], c["scale"], c["inv_scale"],
                                                   d["inv_sigma2"], c["kf_index"], c["px_ref"], c["octave"], c["world_pos"], c["px_curr"])
        return [px, lv, ok], {"tcr": np.stack(tcr)}
    if k == "sares":
        c = d["c"]
        H, J, chi2, nm, vis, patch, jac = R.sparse_align_residuals(c["ref_levels"], c["cur_levels"], d["level"], c["inv_scale"], c["focal"],
                                                                   c["cur_intr"], c["ref_pose"], c["t_cur_ref"], c["kp_xy"], c["valid"], c["world_pos"])
        return [H, J, np.float32(chi2), np.int64(nm), vis, patch, jac], {}
    if k == "sarun":
        c = d["c"]
        t0, T, ret, st = R.sparse_align_run(c["ref_levels"], c["cur_levels"], d["max_level"], d["min_level"], c["inv_scale"], c["focal"],
                                            c["cur_intr"], c["ref_pose"], d["cur_pose"], c["kp_xy"], c["valid"], c["world_pos"])
        return [T, np.int64(ret), np.float32(st[0])], {"t0": t0}
    raise KeyError(k)


def apply_derived(d, der):
    """Puts the reference-derived inputs of a case in place (what the restatement and the kernels are given)."""
    k = d["kind"]
    if k == "kf" and "pose15" in der:
        d["pts"] = dict(d["pts"], pose=np.concatenate([der["pose15"], d["pts"]["pose"][15:]]).astype(np.float32))
    if k == "sim3":
        d["derived"] = np.asarray(der["derived"], np.float32)
    if k == "fdp":
        for kf, t in zip(d["c"]["kfs"], der["tcr"]):
            kf["t_cur_ref"] = np.asarray(t, np.float32)
    if k == "sarun":
        d["c"]["t_cur_ref"] = np.asarray(der["t0"], np.float32)   # cur->mTcw * ref->mTcw.inverse() as the reference composed it
    return d


def sim3_directions(d):
    """SearchBySim3 as two one-direction searches: (pts1 with pose R1w, t1w, sR21, t21), (pts2 with R2w, t2w, sR12, t12); the
    points vpMatches12 holds on entry are switched off on both sides (src/ORBmatcher.cc:1033-1041)."""
    c, der = d["c"], d["derived"]
    n = len(d["kps"])
    am1 = c["already12"] < 0
    am2 = np.ones(n, bool)
    am2[c["already12"][c["already12"] >= 0]] = False

    def one(pts, pose_w, A, b, keep):
        pose = np.zeros(30, np.float32)
        pose[:12] = pose_w; pose[12:21] = A; pose[21:24] = b; pose[24:28] = c["intr"]
        return dict(pts, pose=pose, valid=(pts["valid"] & keep).astype(np.uint8))
    return one(c["pts1"], c["pose1"], der[9:18], der[18:21], am1.astype(np.uint8)), one(c["pts2"], c["pose2"], der[0:9], c["t12"], am2.astype(np.uint8))


# ---- the restatement (oracle/match_oracle.cpp, direct_oracle.cpp) -------------------------------------------------------------
def run_oracle(d):
    sf, isf, s2, inv_s2 = tables()
    k = d["kind"]
    if k == "sbp":
        nm, a, _ = O.search_by_projection(d["kps"], d["desc"], d["u_right"], d["claimed"], BOUNDS, sf, d["mp"], d["th"], d["check_level"], d["ratio"])
        return [np.int32(nm), a]
    if k == "last":
        nm, a, _ = O.search_by_projection_last(d["kps"], d["desc"], d["u_right"], d["claimed"], BOUNDS, sf, d["last"], d["th"], d["mono"],
                                               d["check_level"], d["check_ori"])
        return [np.int32(nm), a]
    if k == "kf":
        nm, a, b = O.search_by_projection_keyframe(d["kps"], d["desc"], d["claimed"], BOUNDS, sf, d["pts"], d["th"], d["variant"],
                                                   d["orb_dist"], d["check_ori"], d["u_right"], d["inv_sigma2"])
        return [np.int32(nm), a] + ([b] if d["variant"] >= 2 else [])
    if k == "sim3":
        p1, p2 = sim3_directions(d)
        m1 = O.search_by_projection_keyframe(d["kps"], d["desc"], None, BOUNDS, sf, p1, d["th"], 4, 100)[2]
        m2 = O.search_by_projection_keyframe(d["kps"], d["desc"], None, BOUNDS, sf, p2, d["th"], 4, 100)[2]
        nf, m12 = O.sim3_agreement(m1, m2)
        return [np.int32(nf), np.where(d["c"]["already12"] >= 0, d["c"]["already12"], m12).astype(np.int32)]
    if k == "init":
        nm, m12, prev = O.search_for_initialization(d["k1"], d["d1"], d["k2"], d["d2"], BOUNDS, d["prev"], d["window"], d["ratio"], d["check_ori"])
        return [np.int32(nm), m12, prev]
    if k == "tri":
        nm, m = O.search_for_triangulation(d["kf1"], d["kf2"], d["F12"], d["ex"], d["ey"], d["only_stereo"], d["check_ori"])
        return [np.int32(nm), m]
    if k == "bow":
        nm, m = O.search_by_bow(d["kf1"], d["kf2"], d["frame2"], d["ratio"], d["check_ori"])
        return [np.int32(nm), m]
    if k == "stereo":
        ur, dp, _ = O.stereo_match(d["kl"], d["dl"], d["kr"], d["dr"], d["pyr_l"], d["pyr_r"], sf, isf, d["bf"], d["b"])
        return [ur, dp]
    if k == "distinct":
        idx, med, out = O.distinctive_descriptors(d["obs_desc"], d["obs_start"])
        ch = (idx >= 0).astype(np.uint8)
        return [ch, out * ch[:, None]]
    if k == "fdp":
        px, lv, ok = O.find_direct_projection(**d["c"])
        return [px, lv, ok]
    if k == "sares":
        c = d["c"]
        H, J, chi2, nm, vis, patch, jac = O.sparse_align_residuals(c["ref_levels"], c["cur_levels"], d["level"], c["inv_scale"][d["level"]], c["focal"],
                                                                   c["cur_intr"], c["ref_pose"], c["t_cur_ref"], c["kp_xy"], c["valid"], c["world_pos"],
                                                                   accumulate=0)
        mean = np.float32(chi2) / np.float32(nm) if nm else np.float32(np.nan)   # the reference returns chi2 / n_meas_ (:218)
        return [H, J, np.float32(mean), np.int64(nm), vis, patch, jac]
    if k == "sarun":
        c = d["c"]
        T, ret, st = O.sparse_align_run(c["ref_levels"], c["cur_levels"], d["max_level"], d["min_level"], c["inv_scale"], c["focal"], c["cur_intr"],
                                        c["ref_pose"], c["kp_xy"], c["valid"], c["world_pos"], c["t_cur_ref"], accumulate=0)
        return [T, np.int64(ret), np.float32(st[1])]
    raise KeyError(k)


# ---- the kernels through the C ABI -----------------------------------------------------------------------------------------------
def run_gpu(Y, matcher, d):
    sf, isf, s2, inv_s2 = tables()
    k = d["kind"]
    if k == "sbp":
        g = matcher.search_by_projection([(d["kps"], d["desc"], d["u_right"], d["claimed"])], [d["mp"]], BOUNDS, sf, d["th"], d["check_level"], d["ratio"])[0]
        return [np.int32(g[0]), g[1]]
    if k == "last":
        g = matcher.search_by_projection_last([(d["kps"], d["desc"], d["u_right"], d["claimed"])], [d["last"]], BOUNDS, sf, d["th"], d["mono"],
                                              d["check_level"], d["check_ori"])[0]
        return [np.int32(g[0]), g[1]]
    if k == "kf":
        g = matcher.search_by_projection_keyframe([(d["kps"], d["desc"], d["claimed"], d["u_right"])], [d["pts"]], BOUNDS, sf, d["th"],
                                                  d["variant"], d["orb_dist"], d["check_ori"], d["inv_sigma2"])[0]
        if d["variant"] >= 2:   # Fuse: the ABI returns the keypoint per point; the per-keypoint view = last point that asked for it
            kp_to_pt = np.full(len(d["kps"]), -1, np.int32)
            for i, idx in enumerate(g[2]):
                if idx >= 0:
                    kp_to_pt[idx] = i
            return [np.int32(g[0]), kp_to_pt, g[2]]
        return [np.int32(g[0]), g[1]]
    if k == "sim3":
        p1, p2 = sim3_directions(d)
        g = matcher.search_by_sim3([(d["kps"], d["desc"])], [(d["kps"], d["desc"])], [p1], [p2], BOUNDS, BOUNDS, sf, d["th"])[0]
        return [np.int32(g[0]), np.where(d["c"]["already12"] >= 0, d["c"]["already12"], g[1]).astype(np.int32)]
    if k == "init":
        g = matcher.search_for_initialization([(d["k1"], d["d1"])], [(d["k2"], d["d2"])], [d["prev"]], BOUNDS, d["window"], d["ratio"], d["check_ori"])[0]
        return [np.int32(g[0]), g[1], g[2]]
    if k == "tri":
        g = matcher.search_for_triangulation(d["kf1"], d["kf2"], d["F12"], d["ex"], d["ey"], d["only_stereo"], d["check_ori"])
        return [np.int32(g[0]), g[1]]
    if k == "bow":
        g = matcher.search_by_bow(d["kf1"], d["kf2"], d["frame2"], d["ratio"], d["check_ori"])
        return [np.int32(g[0]), g[1]]
    if k == "stereo":
        exl, exr = Y.Extractor(d["nf"]), Y.Extractor(d["nf"])
        kl, dl = exl.extract(d["left"])
        kr, dr = exr.extract(d["right"])
        assert kl.tobytes() == d["kl"].tobytes() and kr.tobytes() == d["kr"].tobytes()
        ur, dp = matcher.stereo_match(exl, exr, 0, kl, dl, kr, dr, d["bf"], d["b"])
        exl.close(); exr.close()
        return [ur, dp]
    if k == "distinct":
        idx, med, out = matcher.distinctive_descriptors(d["obs_desc"], d["obs_start"])
        ch = (idx >= 0).astype(np.uint8)
        return [ch, out * ch[:, None]]
    if k == "fdp":
        c = d["c"]
        pyrs = [Y.Pyramid(kf["levels"]) for kf in c["kfs"]]
        cur = Y.Pyramid(c["cur_levels"])
        kfs = [dict(kf, pyramid=p) for kf, p in zip(c["kfs"], pyrs)]
        args = {a: v for a, v in c.items() if a not in ("kfs", "cur_levels")}
        px, lv, ok = matcher.find_direct_projection(kfs, cur, **args)
        for p in pyrs + [cur]:
            p.close()
        return [px, lv, ok]
    raise KeyError(k)


def same(a, b):
    """Outputs equal: integers by value, floats by their bytes (NaN == NaN)."""
    if len(a) != len(b):
        return False
    for x, y in zip(a, b):
        x, y = np.asarray(x), np.asarray(y)
        if x.shape != y.shape or x.dtype != y.dtype or x.tobytes() != y.tobytes():
            return False
    return True
