"""The drop-in C++ classes (orb-ygz-slam_b200/host: ygz::ORBextractor, ygz::ORBmatcher) driven the way
the reference's callers drive them (host_demo.cc), compared with the oracle."""
import os
import subprocess
import tempfile

import numpy as np
import pytest

import oracle_lib as O
import synth
from ygz import Y, PKG

pytestmark = pytest.mark.gpu
DEMO = os.path.join(PKG, "host", "host_demo")


def test_host_classes_end_to_end():
    if not os.path.exists(DEMO):
        subprocess.run(["make", "-C", os.path.join(PKG, "host")], check=True)
    w, h, nf, n_mp = 752, 480, 1000, 3000
    left, right = synth.stereo_pair(w, h, 11)
    p = O.params(nf)
    kl, dl = O.extract(p, left)
    kr, dr = O.extract(p, right)
    mp, _, _ = synth.projection_case(kl, dl, n_mp, 3)
    mp["in_view"][:] = 1; mp["has_obs"][:] = 1
    with tempfile.TemporaryDirectory() as td:
        fin, fout = os.path.join(td, "in.bin"), os.path.join(td, "out.bin")
        with open(fin, "wb") as f:
            np.array([w, h, nf, n_mp], np.int32).tofile(f)
            left.tofile(f); right.tofile(f)
            rec = np.zeros(n_mp, np.dtype([("v", "<f4", 3), ("lvl", "<i4"), ("d", "u1", 32)]))
            rec["v"][:, 0] = mp["proj_x"]; rec["v"][:, 1] = mp["proj_y"]; rec["v"][:, 2] = mp["view_cos"]
            rec["lvl"] = mp["level"]; rec["d"] = mp["desc"]
            rec.tofile(f)
            n_last = 900
            last, _, _ = synth.lastframe_case(kl, dl, n_last, 21)
            np.array([n_last], np.int32).tofile(f)
            last["pose"].tofile(f)
            lrec = np.zeros(n_last, np.dtype([("v", "<f4", 4), ("q", "<i4", 3), ("d", "u1", 32)]))
            lrec["v"][:, :3] = last["world_pos"]; lrec["v"][:, 3] = last["angle"]
            lrec["q"][:, 0] = last["octave"]; lrec["q"][:, 1] = last["valid"]; lrec["q"][:, 2] = last["has_obs"]
            lrec["d"] = last["desc"]
            lrec.tofile(f)
        # a vocabulary file in the reference's binary format: the golden one its own DBoW2 wrote (k=10, L=3)
        g = np.load(os.path.join(O.ROOT, "tests", "golden", "bow_k10L3.npz"))
        fvoc, fbow = os.path.join(td, "voc.bin"), os.path.join(td, "bow.bin")
        open(fvoc, "wb").write(g["voc_bin"].tobytes())
        res = subprocess.run([DEMO, fin, fout, fvoc, fbow], capture_output=True, text=True)
        assert res.returncode == 0, res.stdout + res.stderr
        buf = open(fout, "rb").read()
        bowbuf = open(fbow, "rb").read()
        ovoc = O.Vocabulary.load_binary(fvoc)
    off = 0

    def take(dtype, n):
        nonlocal off
        a = np.frombuffer(buf, dtype, n, off)
        off += a.nbytes
        return a
    nL = int(take("<i4", 1)[0]); gkl = take(Y.KP_DTYPE, nL); gdl = take(np.uint8, nL * 32).reshape(nL, 32)
    nR = int(take("<i4", 1)[0]); gkr = take(Y.KP_DTYPE, nR); gdr = take(np.uint8, nR * 32).reshape(nR, 32)
    gur = take("<f4", nL); gdepth = take("<f4", nL)
    gnm = int(take("<i4", 1)[0]); gassign = take("<i4", nL)
    assert gkl.tobytes() == kl.tobytes() and np.array_equal(gdl, dl)
    assert gkr.tobytes() == kr.tobytes() and np.array_equal(gdr, dr)
    sf, isf, _ = synth.scale_tables()
    bf = np.float32(386.1448)
    ur, depth, _ = O.stereo_match(kl, dl, kr, dr, O.pyramid(p, left), O.pyramid(p, right), sf, isf, float(bf), float(bf / np.float32(718.856)))
    assert gur.tobytes() == ur.tobytes() and gdepth.tobytes() == depth.tobytes() and (ur >= 0).sum() > 100
    nm, kp_to_mp, _ = O.search_by_projection(kl, dl, None, None, (0.0, 0.0, float(w), float(h)), sf, mp, 1.0, False, 0.8)
    assert gnm == nm and np.array_equal(gassign, kp_to_mp) and nm > 300
    gnml = int(take("<i4", 1)[0]); gassign_last = take("<i4", nL)
    nml, kp_to_pt, _ = O.search_by_projection_last(kl, dl, None, None, (0.0, 0.0, float(w), float(h)), sf, last, 15.0, True, True, True)
    assert gnml == nml and np.array_equal(gassign_last, kp_to_pt) and nml > 300
    # DSO_KEYPOINT through the host class == the reference's own code on the same stream of two frames
    import refcli
    if refcli.available():
        exist = kl[::7].copy()
        want_dso = refcli.run_ref_dso(np.stack([left, left]), [exist, exist], nfeatures=nf)
        for k in range(2):
            nd = int(take("<i4", 1)[0]); gk = take(Y.KP_DTYPE, nd); gd = take(np.uint8, nd * 32).reshape(nd, 32)
            assert gk.tobytes() == want_dso[k]["kps"].tobytes() and np.array_equal(gd, want_dso[k]["desc"]) and nd > len(exist) + 300
        # the right eye of a stereo frame that came out of direct tracking (Frame N > 0): plain OctTree on mImRight
        # (src/ORBextractor.cc:1037-1038, 1059-1060), and the Right extractor's pyramid is the right image's
        nr = int(take("<i4", 1)[0]); gk = take(Y.KP_DTYPE, nr); gd = take(np.uint8, nr * 32).reshape(nr, 32)
        assert gk.tobytes() == kr.tobytes() and np.array_equal(gd, dr) and nr > 900
        for lv in O.pyramid(p, right):
            lw, lh = take("<i4", 2)
            assert (lh, lw) == lv.shape and np.array_equal(take(np.uint8, int(lw) * int(lh)).reshape(lv.shape), lv)
        # ORBSLAM_KEYPOINT on a left Frame that already holds keypoints == the reference's own code (:1086-1099)
        want_e = refcli.run_ref_dso(left[None], [exist], nfeatures=nf, orbslam=True)[0]
        ne = int(take("<i4", 1)[0]); gk = take(Y.KP_DTYPE, ne); gd = take(np.uint8, ne * 32).reshape(ne, 32)
        assert ne == len(exist) + len(kl) and gk.tobytes() == want_e["kps"].tobytes() and np.array_equal(gd, want_e["desc"])
    # ORBVocabulary::transform through the host class == the oracle (pinned to DBoW2) on the same descriptors
    want = ovoc.transform(dl, 4)
    nb = int(np.frombuffer(bowbuf, "<i4", 1, 0)[0])
    b = np.frombuffer(bowbuf, np.dtype([("w", "<i4"), ("v", "<f8")]), nb, 4)
    assert np.array_equal(b["w"], want["bow_word"]) and b["v"].tobytes() == want["bow_value"].tobytes() and nb > 100
    o = 4 + nb * 12
    nn = int(np.frombuffer(bowbuf, "<i4", 1, o)[0]); o += 4
    assert nn == len(want["fv_node"])
    for j in range(nn):
        node, cnt = np.frombuffer(bowbuf, "<i4", 2, o); o += 8
        feats = np.frombuffer(bowbuf, "<i4", int(cnt), o); o += 4 * int(cnt)
        assert node == want["fv_node"][j] and np.array_equal(feats, want["fv_feat"][want["fv_start"][j]:want["fv_start"][j + 1]])
    # ORBmatcher::SearchByBoW, both overloads, on keyframes made of the two images (MapPoint pattern of host_demo.cc)
    wantR = ovoc.transform(dr, 4)
    def side(k, d, r):
        i = np.arange(len(k))
        return {"kps": k, "desc": d, "has_mp": ((i % 3 != 1) & (i % 7 != 3)).astype(np.uint8), "node_ids": r["fv_node"],
                "node_start": r["fv_start"], "node_feats": r["fv_feat"]}
    gn = int(np.frombuffer(bowbuf, "<i4", 1, o)[0]); gm = np.frombuffer(bowbuf, "<i4", nL, o + 4); o += 4 + 4 * nL
    en, em = O.search_by_bow(side(kr, dr, wantR), side(kl, dl, want), True, 0.7, True)
    assert gn == en and np.array_equal(gm, em) and en > 100
    gn = int(np.frombuffer(bowbuf, "<i4", 1, o)[0]); gm = np.frombuffer(bowbuf, "<i4", nL, o + 4); o += 4 + 4 * nL
    en, em = O.search_by_bow(side(kl, dl, want), side(kr, dr, wantR), False, 0.75, True)
    assert gn == en and np.array_equal(gm, em) and en > 50
    assert o == len(bowbuf)
    ovoc.close()


DEMO2 = os.path.join(PKG, "host", "host_demo2")
REC = np.dtype([("wpos", "<f4", 3), ("normal", "<f4", 3), ("angle", "<f4"), ("max_d", "<f4"), ("min_d", "<f4"),
                ("valid", "<i4"), ("n_obs", "<i4"), ("desc", "u1", 32)])


def _records(pts, n_obs):
    r = np.zeros(len(pts["valid"]), REC)
    r["wpos"] = pts["world_pos"]; r["normal"] = pts["normal"]; r["angle"] = pts["angle"]
    r["max_d"] = pts["max_distance"]; r["min_d"] = pts["min_distance"]; r["valid"] = pts["valid"]; r["n_obs"] = n_obs
    r["desc"] = pts["desc"]
    return r


def test_host_matcher_overloads_on_objects():
    """SearchForInitialization, the relocalisation SearchByProjection and Fuse through ygz::ORBmatcher on Frame / KeyFrame /
    MapPoint objects (host_demo2.cc), compared with the oracle; Fuse's map updates are replayed here from the oracle's
    per-point result with the rules of src/ORBmatcher.cc:862-877."""
    if not os.path.exists(DEMO2):
        subprocess.run(["make", "-C", os.path.join(PKG, "host")], check=True)
    w, h, nf = 752, 480, 1000
    img = synth.textured_frame(w, h, 5)
    kps, desc = O.extract(O.params(nf), img)
    n = len(kps)
    sf, _, sigma2 = synth.scale_tables()
    inv_sigma2 = (np.float32(1.0) / sigma2).astype(np.float32)
    bounds = (0.0, 0.0, float(w), float(h))
    rng = np.random.default_rng(9)
    _, _, k2, d2, _ = synth.initialization_case(kps, desc, 31)
    prev = np.stack([kps["x"], kps["y"]], 1).astype(np.float32)
    reloc, claimed = synth.keyframe_points_case(kps, desc, 900, 32, 0)
    fuse, _ = synth.keyframe_points_case(kps, desc, 1100, 33, 2)
    fuse_obs = rng.integers(0, 6, 1100).astype(np.int32)
    existing = np.where(rng.uniform(size=n) < 0.4, rng.integers(0, 6, n), -1).astype(np.int32)
    with tempfile.TemporaryDirectory() as td:
        fin, fout = os.path.join(td, "in.bin"), os.path.join(td, "out.bin")
        with open(fin, "wb") as f:
            np.array([n], np.int32).tofile(f); kps.tofile(f); desc.tofile(f)
            np.array(bounds, np.float32).tofile(f); np.array([len(sf)], np.int32).tofile(f); sf.tofile(f); inv_sigma2.tofile(f)
            np.array([reloc["log_scale_factor"]], np.float32).tofile(f)
            np.array([len(k2)], np.int32).tofile(f); k2.tofile(f); d2.tofile(f)
            np.array([100], np.int32).tofile(f); np.array([0.9], np.float32).tofile(f)
            np.array([len(reloc["valid"])], np.int32).tofile(f); reloc["pose"].tofile(f); _records(reloc, 1).tofile(f)
            claimed.tofile(f); np.array([10.0], np.float32).tofile(f); np.array([100], np.int32).tofile(f)
            np.array([len(fuse["valid"])], np.int32).tofile(f); fuse["pose"].tofile(f); _records(fuse, fuse_obs).tofile(f)
            fuse["u_right"].tofile(f); existing.tofile(f); np.array([3.0], np.float32).tofile(f)
        res = subprocess.run([DEMO2, fin, fout], capture_output=True, text=True)
        assert res.returncode == 0, res.stdout + res.stderr
        buf = open(fout, "rb").read()
    off = 0

    def take(dtype, cnt):
        nonlocal off
        a = np.frombuffer(buf, dtype, cnt, off)
        off += a.nbytes
        return a
    # SearchForInitialization
    en, em, ep = O.search_for_initialization(kps, desc, k2, d2, bounds, prev, 100, float(np.float32(0.9)), True)
    assert int(take("<i4", 1)[0]) == en and en > 50
    assert np.array_equal(take("<i4", n), em)
    got_prev = take("<f4", 2 * n).reshape(n, 2)
    assert got_prev.tobytes() == ep.tobytes()
    # relocalisation SearchByProjection
    en, e_kp, _ = O.search_by_projection_keyframe(kps, desc, claimed, bounds, sf, reloc, 10.0, 0, 100, True)
    assert int(take("<i4", 1)[0]) == en and en > 100
    want = np.where(e_kp >= 0, e_kp, np.where(claimed > 0, -2, -1))
    assert np.array_equal(take("<i4", n), want)
    # Fuse: replay :862-877 over the oracle's per-point keypoints
    m = len(fuse["valid"])
    _, _, best = O.search_by_projection_keyframe(kps, desc, None, bounds, sf, fuse, 3.0, 2, 50, False, fuse["u_right"], inv_sigma2)
    slot = np.where(existing >= 0, m + np.arange(n), -1)
    obs = np.concatenate([fuse_obs, np.where(existing >= 0, existing, 1)]).astype(np.int64)   # MapPointLite default nObs = 1
    bad = np.concatenate([fuse["valid"] == 0, np.zeros(n, bool)])
    in_kf = np.concatenate([np.zeros(m, bool), existing >= 0])
    kf_idx = np.concatenate([np.full(m, -1), np.arange(n)])
    replaced = np.full(m + n, -1)
    n_fused = 0

    def replace(a, b):          # a->Replace(b), MapPointLite::Replace
        bad[a] = True; replaced[a] = b
        if in_kf[a]:
            in_kf[a] = False
            if not in_kf[b]:
                slot[kf_idx[a]] = b; in_kf[b] = True; kf_idx[b] = kf_idx[a]; obs[b] += 1
            else:
                slot[kf_idx[a]] = -1
    for i in range(m):
        if best[i] < 0 or bad[i] or in_kf[i]:
            continue
        cur = slot[best[i]]
        if cur >= 0:
            if not bad[cur]:
                if obs[cur] > obs[i]:
                    replace(i, cur)
                else:
                    replace(cur, i)
        else:
            slot[best[i]] = i; in_kf[i] = True; kf_idx[i] = best[i]; obs[i] += 1
        n_fused += 1
    assert int(take("<i4", 1)[0]) == n_fused and n_fused > 100
    assert np.array_equal(take("<i4", n), slot)
    assert np.array_equal(take("<i4", m + n), replaced)
    assert np.array_equal(take("<i4", m + n), obs)
    assert off == len(buf)


DEMO3 = os.path.join(PKG, "host", "host_demo3")


def test_host_direct_tracking_on_objects():
    """SparseImgAlign::run, ORBmatcher::FindDirectProjection (batched) and MapPoint::ComputeDistinctiveDescriptors through the
    host drop-ins of host/DirectTracking.h on Frame / KeyFrame / MapPoint objects (host_demo3.cc) against the oracle.  The host
    composes T_cur_from_ref with its pose type's own product, so floats agree to rounding of that product, not bit for bit."""
    from test_direct_oracle import synth_observations
    if not os.path.exists(DEMO3):
        subprocess.run(["make", "-C", os.path.join(PKG, "host")], check=True)
    c = synth.sparse_align_case(4, n=700)
    n = len(c["valid"])
    rng = np.random.default_rng(3)
    valid = c["valid"].copy()
    bad = np.where(valid == 1)[0]
    valid[bad[:20]] = 2; valid[bad[20:40]] = 3                      # outliers / bad MapPoints: not used (src/SparseImageAlign.cc:68-70)
    used = (valid == 1).astype(np.uint8)
    kps = np.zeros(n, O.KP_DTYPE)
    kps["x"] = c["kp_xy"][:, 0]; kps["y"] = c["kp_xy"][:, 1]; kps["octave"] = rng.integers(0, 4, n); kps["size"] = 31; kps["angle"] = -1
    sc, inv, sigma2 = synth.scale_tables()
    inv_sigma2 = (np.float32(1.0) / sigma2).astype(np.float32)
    # cur->mTcw such that cur->mTcw * ref->mTcw.inverse() is the case's initial guess: cur = guess * ref
    q_g, t_g, q_r, t_r = c["t_cur_ref"][:4], c["t_cur_ref"][4:], c["ref_pose"][:4], c["ref_pose"][4:]

    def qmul(a, b):
        ax, ay, az, aw = (float(v) for v in a); bx, by, bz, bw = (float(v) for v in b)
        return np.array([aw * bx + ax * bw + ay * bz - az * by, aw * by + ay * bw + az * bx - ax * bz, aw * bz + az * bw + ax * by - ay * bx,
                         aw * bw - ax * bx - ay * by - az * bz])
    q_c = qmul(q_g, q_r); q_c /= np.linalg.norm(q_c)
    t_c = synth.quat_rotate(q_g, t_r) + t_g.astype(np.float64)
    cur_pose = np.concatenate([q_c, t_c]).astype(np.float32)
    # direct-projection requests on valid interior features
    sel = np.where((used == 1) & (c["kp_xy"][:, 0] > 40) & (c["kp_xy"][:, 0] < 700) & (c["kp_xy"][:, 1] > 40) & (c["kp_xy"][:, 1] < 440))[0][:200]
    fx, fy, cx, cy = (float(v) for v in c["cur_intr"])
    px0 = np.zeros((len(sel), 2), np.float32)
    t_true_pose = np.concatenate([[0, 0, 0, 1], c["t_true"]]).astype(np.float32)
    for j, i in enumerate(sel):
        p_ref = synth.quat_rotate(q_r, c["world_pos"][i]) + t_r.astype(np.float64)
        p_cur = p_ref + c["t_true"]
        px0[j] = [fx * p_cur[0] / p_cur[2] + cx, fy * p_cur[1] / p_cur[2] + cy]
    px0 += rng.uniform(-1.2, 1.2, px0.shape).astype(np.float32)
    desc, start = synth_observations(50, 7)
    with tempfile.TemporaryDirectory() as td:
        fin, fout = os.path.join(td, "in.bin"), os.path.join(td, "out.bin")
        with open(fin, "wb") as f:
            np.array([8], np.int32).tofile(f)
            for pyr in (c["ref_levels"], c["cur_levels"]):
                for im in pyr:
                    np.array([im.shape[1], im.shape[0]], np.int32).tofile(f); np.ascontiguousarray(im).tofile(f)
            sc.tofile(f); inv.tofile(f); inv_sigma2.tofile(f); c["cur_intr"].tofile(f)
            c["ref_pose"].astype(np.float32).tofile(f); cur_pose.tofile(f)
            np.array([n], np.int32).tofile(f); kps.tofile(f); valid.tofile(f); c["world_pos"].tofile(f)
            np.array([len(sel)], np.int32).tofile(f)
            for j, i in enumerate(sel):
                np.array([i], np.int32).tofile(f); px0[j].tofile(f)
            np.array([len(start) - 1], np.int32).tofile(f); start.tofile(f); desc.tofile(f)
        res = subprocess.run([DEMO3, fin, fout], capture_output=True, text=True)
        assert res.returncode == 0, res.stdout + res.stderr
        with open(fout, "rb") as f:
            T = np.fromfile(f, np.float32, 7); tracked = int(np.fromfile(f, np.int64, 1)[0])
            dp = np.fromfile(f, np.dtype([("px", "<f4", 2), ("level", "<i4"), ("ok", "<i4")]), len(sel))
            best = np.fromfile(f, np.uint8, (len(start) - 1) * 32).reshape(-1, 32)
    # SparseImgAlign::run
    args = {k: v for k, v in c.items() if k not in ("t_true", "valid")}
    eT, en, _ = O.sparse_align_run(max_level=7, min_level=1, accumulate=1, valid=used, **args)
    assert tracked == en and np.abs(T - eT).max() < 2e-4 and np.abs(T[4:] - c["t_true"]).max() < 2e-3
    # FindDirectProjection: the keyframe is the reference frame; TCR = the host's product of the two poses ~ the initial guess
    kf = dict(levels=c["ref_levels"], pose_ref=c["ref_pose"], t_cur_ref=c["t_cur_ref"], fx=fx, fy=fy, cx=cx, cy=cy)
    e_px, e_lv, e_ok = O.find_direct_projection([kf], c["cur_levels"], c["cur_intr"], sc, inv, float(inv_sigma2[1]),
                                                np.zeros(len(sel), np.int32), c["kp_xy"][sel], kps["octave"][sel], c["world_pos"][sel], px0)
    assert np.array_equal(dp["level"], e_lv)
    assert (dp["ok"] == e_ok).mean() > 0.97 and e_ok.mean() > 0.5
    both = (dp["ok"] == 1) & (e_ok == 1)
    assert np.abs(dp["px"][both] - e_px[both]).max() < 5e-3
    # ComputeDistinctiveDescriptors
    marker = np.full((len(start) - 1, 32), 0xA5, np.uint8)
    _, _, e_best = O.distinctive_descriptors(desc, start, marker)
    assert np.array_equal(best, e_best)
