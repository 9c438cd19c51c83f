"""GPU parity of the Hamming matching paths (libygzorb_cuda.so through the C ABI) against the CPU
oracle (oracle/match_oracle.cpp): indices and integer distances bit-exact; stereo uRight / depth
floats compared bit-exact (same float op order, no contraction)."""
import numpy as np
import pytest

import oracle_lib as O
import synth
from ygz import Y

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def matcher():
    m = Y.Matcher()
    yield m
    m.close()


@pytest.fixture(scope="module")
def feats():
    img = synth.textured_frame(752, 480, 0)
    ex = Y.Extractor(1000)
    kps, desc = ex.extract(img)
    ex.close()
    return img, kps, desc


def test_descriptor_distance(matcher):
    rng = np.random.default_rng(0)
    a = rng.integers(0, 256, size=(1000, 32), dtype=np.uint8)
    b = rng.integers(0, 256, size=(1000, 32), dtype=np.uint8)
    b[:10] = a[:10]; b[10] = ~a[10]
    d = matcher.descriptor_distance(a, b)
    exp = np.unpackbits(a ^ b, axis=1).sum(axis=1)
    assert np.array_equal(d, exp) and d[0] == 0 and d[10] == 256


@pytest.mark.parametrize("variant", [0, 1, 2, 3])
@pytest.mark.parametrize("nq,nt,ratio,th", [(1000, 1500, 0.6, 50), (3, 1, 0.9, 100), (513, 512, 0.75, 256), (700, 70001, 0.9, 100),
                                            (5, 0, 0.6, 50), (2049, 513, 0.6, 50)])
def test_knn2_vs_oracle(matcher, variant, nq, nt, ratio, th):
    q, t = synth.knn_case(nq, max(nt, 1), nq + nt)
    t = t[:nt]
    if nt > 10:
        t[7] = t[3]; q[0] = t[3]          # exact duplicate rows: the first index must win, d2 == d1 == 0
    matcher.set_knn_variant(variant)
    a = matcher.hamming_knn2(q, t, ratio, th)
    b = O.hamming_knn2(q, t, ratio, th, threads=8)
    for x, y, name in zip(a, b, ("idx", "d1", "d2", "match")):
        assert np.array_equal(x, y), name
    matcher.set_knn_variant(2)


@pytest.mark.parametrize("th,check_level,stereo,seed", [(1.0, False, False, 1), (5.0, False, True, 2), (3.0, True, False, 3),
                                                         (15.0, False, False, 4)])
def test_search_by_projection_vs_oracle(matcher, feats, th, check_level, stereo, seed):
    img, kps, desc = feats
    sf, _, _ = synth.scale_tables()
    bounds = (0.0, 0.0, float(img.shape[1]), float(img.shape[0]))
    frames, mps, exp = [], [], []
    for k in range(3):                                     # a small batch, ragged MapPoint counts
        mp, u_right, claimed = synth.projection_case(kps, desc, 3000 - 500 * k, seed * 10 + k, stereo=stereo)
        frames.append((kps, desc, u_right, claimed)); mps.append(mp)
        exp.append(O.search_by_projection(kps, desc, u_right, claimed, bounds, sf, mp, th, check_level, 0.8))
    got = matcher.search_by_projection(frames, mps, bounds, sf, th, check_level, 0.8)
    for g, e in zip(got, exp):
        assert g[0] == e[0] and e[0] > 100
        assert np.array_equal(g[1], e[1]) and np.array_equal(g[2], e[2])


@pytest.mark.parametrize("th,mono,check_level,check_ori,stereo,motion,seed", [
    (15.0, True, False, True, False, (0.02, -0.01, 0.3), 1), (7.0, False, True, True, True, (0.0, 0.0, 0.5), 2),
    (7.0, False, True, False, True, (0.0, 0.0, -0.5), 3), (15.0, False, True, True, False, (0.05, 0.0, 0.01), 4),
    (40.0, True, False, True, False, (0.0, 0.0, 0.0), 5)])
def test_search_by_projection_last_vs_oracle(matcher, feats, th, mono, check_level, check_ori, stereo, motion, seed):
    """SearchByProjection(CurrentFrame, LastFrame, ...) (src/ORBmatcher.cc:1218-1350): forward / backward / lateral
    motion (octave windows), stereo consistency, claim order, rotation-histogram rejection; th=40 overflows the
    candidate cache."""
    img, kps, desc = feats
    sf, _, _ = synth.scale_tables()
    bounds = (0.0, 0.0, float(img.shape[1]), float(img.shape[0]))
    frames, lasts, exp = [], [], []
    for k in range(3):                                     # a small batch, ragged point counts
        last, u_right, claimed = synth.lastframe_case(kps, desc, 1000 - 150 * k, seed * 10 + k, motion=motion, stereo=stereo)
        frames.append((kps, desc, u_right, claimed)); lasts.append(last)
        exp.append(O.search_by_projection_last(kps, desc, u_right, claimed, bounds, sf, last, th, mono, check_level, check_ori))
    got = matcher.search_by_projection_last(frames, lasts, bounds, sf, th, mono, check_level, check_ori)
    for g, e in zip(got, exp):
        assert g[0] == e[0] and e[0] > 100
        assert np.array_equal(g[1], e[1]) and np.array_equal(g[2], e[2])


@pytest.mark.parametrize("th,variant,orb_dist,check_ori,seed", [
    (10.0, 0, 100, True, 1), (3.0, 0, 64, True, 2), (15.0, 0, 100, False, 3), (10.0, 1, 50, False, 4),
    (4.0, 1, 50, False, 5), (40.0, 1, 50, False, 6), (3.0, 2, 50, False, 7), (2.5, 2, 50, False, 8), (4.0, 3, 50, False, 9),
    (20.0, 3, 50, False, 10)])
def test_search_by_projection_keyframe_vs_oracle(matcher, feats, th, variant, orb_dist, check_ori, seed):
    """The relocalisation (src/ORBmatcher.cc:1352-1469) and Sim3 (:265-373) projection searches and the search halves of
    Fuse (:748-886) / Fuse with Sim3 (:888-1004): in-kernel projection, distance range, viewing angle, PredictScale,
    level window, chi-square gate, claim order, rotation rejection."""
    img, kps, desc = feats
    sf, _, sigma2 = synth.scale_tables()
    inv_sigma2 = (np.float32(1.0) / sigma2).astype(np.float32)
    bounds = (0.0, 0.0, float(img.shape[1]), float(img.shape[0]))
    frames, pts, exp = [], [], []
    for k in range(3):
        p, claimed = synth.keyframe_points_case(kps, desc, 1200 - 200 * k, seed * 10 + k, variant)
        ur = p["u_right"] if variant == 2 else None
        frames.append((kps, desc, claimed, ur)); pts.append(p)
        exp.append(O.search_by_projection_keyframe(kps, desc, claimed, bounds, sf, p, th, variant, orb_dist, check_ori, ur, inv_sigma2))
    got = matcher.search_by_projection_keyframe(frames, pts, bounds, sf, th, variant, orb_dist, check_ori, inv_sigma2)
    for g, e in zip(got, exp):
        assert g[0] == e[0] and e[0] > 100
        assert (g[1] is None or np.array_equal(g[1], e[1])) and np.array_equal(g[2], e[2])


@pytest.mark.parametrize("th,seed", [(7.5, 1), (3.0, 2)])
def test_search_by_sim3_vs_oracle(matcher, feats, th, seed):
    """SearchBySim3 (src/ORBmatcher.cc:1006-1216): both directions + the agreement check.  The two keyframes share the
    keypoints here (MapPoint i of either side sits near keypoint src[i] of the other), which makes mutual matches common."""
    img, kps, desc = feats
    sf, _, _ = synth.scale_tables()
    bounds = (0.0, 0.0, float(img.shape[1]), float(img.shape[0]))
    n = len(kps)
    kfs1, kfs2, P1, P2, exp = [], [], [], [], []
    for k in range(2):
        p1, _ = synth.keyframe_points_case(kps, desc, n, seed * 10 + k, 4, jitter=1.0)
        p2, _ = synth.keyframe_points_case(kps, desc, n, seed * 10 + k + 5, 4, jitter=1.0)
        # point i of KF1 should land on keypoint i of KF2 and vice versa for agreement: re-aim both at the identity pairing
        for p in (p1, p2):
            order = np.argsort(p["src"], kind="stable")
            first = np.full(n, -1); first[p["src"][order][::-1]] = order[::-1]
            sel = np.where(first >= 0, first, 0)
            for key in ("world_pos", "normal", "desc", "angle", "max_distance", "min_distance"):
                p[key] = p[key][sel]
            p["valid"] = (p["valid"][sel] & (first >= 0)).astype(np.uint8)
        m1 = O.search_by_projection_keyframe(kps, desc, None, bounds, sf, p1, th, 4, 100)[2]
        m2 = O.search_by_projection_keyframe(kps, desc, None, bounds, sf, p2, th, 4, 100)[2]
        exp.append(O.sim3_agreement(m1, m2))
        kfs1.append((kps, desc)); kfs2.append((kps, desc)); P1.append(p1); P2.append(p2)
    got = matcher.search_by_sim3(kfs1, kfs2, P1, P2, bounds, bounds, sf, th)
    for g, e in zip(got, exp):
        assert g[0] == e[0] and e[0] > 50 and np.array_equal(g[1], e[1])


@pytest.mark.parametrize("window,ratio,check_ori,seed", [(100, 0.9, True, 1), (30, 0.9, True, 2), (100, 0.7, False, 3),
                                                         (400, 0.95, True, 4)])
def test_search_for_initialization_vs_oracle(matcher, feats, window, ratio, check_ori, seed):
    """SearchForInitialization (src/ORBmatcher.cc:375-478): level-0 window search with the vMatchedDistance take-over rule,
    rotation rejection and the vbPrevMatched update; competing near-copies in F1 exercise the serial order."""
    img, kps, desc = feats
    bounds = (0.0, 0.0, float(img.shape[1]), float(img.shape[0]))
    f1, f2, prevs, exp = [], [], [], []
    for k in range(3):
        k1, d1, k2, d2, prev = synth.initialization_case(kps[:len(kps) - 100 * k], desc[:len(kps) - 100 * k], seed * 10 + k)
        f1.append((k1, d1)); f2.append((k2, d2)); prevs.append(prev)
        exp.append(O.search_for_initialization(k1, d1, k2, d2, bounds, prev, window, ratio, check_ori))
    got = matcher.search_for_initialization(f1, f2, prevs, bounds, window, ratio, check_ori)
    for g, e in zip(got, exp):
        assert g[0] == e[0] and e[0] > 100
        assert np.array_equal(g[1], e[1]) and g[2].tobytes() == e[2].tobytes()


def test_search_by_projection_claim_chain(matcher, feats):
    """Many MapPoints aimed at the same keypoints: long claim chains must resolve like the serial loop."""
    img, kps, desc = feats
    sf, _, _ = synth.scale_tables()
    rng = np.random.default_rng(5)
    src = np.repeat(rng.integers(0, len(kps), size=40), 25)
    mp = {"proj_x": kps["x"][src].copy(), "proj_y": kps["y"][src].copy(), "proj_xr": np.zeros(len(src), np.float32),
          "view_cos": np.full(len(src), 0.95, np.float32), "level": kps["octave"][src].astype(np.int32),
          "in_view": np.ones(len(src), np.uint8), "has_obs": (rng.uniform(size=len(src)) < 0.8).astype(np.uint8),
          "desc": synth.flip_bits(desc[src], rng.integers(0, 10, size=len(src)), rng)}
    bounds = (0.0, 0.0, float(img.shape[1]), float(img.shape[0]))
    e = O.search_by_projection(kps, desc, None, None, bounds, sf, mp, 8.0, False, 0.9)
    g = matcher.search_by_projection([(kps, desc, None, None)], [mp], bounds, sf, 8.0, False, 0.9)[0]
    assert g[0] == e[0] and np.array_equal(g[1], e[1]) and np.array_equal(g[2], e[2])


@pytest.mark.parametrize("only_stereo,check_ori,seed", [(False, False, 1), (False, True, 2), (True, True, 3)])
def test_triangulation_vs_oracle(matcher, feats, only_stereo, check_ori, seed):
    _, kps, desc = feats
    kf1, kf2, F12, ex, ey = synth.triangulation_case(kps, desc, seed)
    e = O.search_for_triangulation(kf1, kf2, F12, ex, ey, only_stereo, check_ori)
    g = matcher.search_for_triangulation(kf1, kf2, F12, ex, ey, only_stereo, check_ori)
    assert g[0] == e[0] and np.array_equal(g[1], e[1]) and e[0] > (5 if only_stereo else 20)


@pytest.mark.parametrize("frame2,check_ori,ratio,n_nodes,seed", [(True, True, 0.7, 100, 1), (True, False, 0.9, 30, 2),
                                                                (False, True, 0.75, 100, 3), (False, False, 0.6, 5, 4)])
def test_search_by_bow_vs_oracle(matcher, feats, frame2, check_ori, ratio, n_nodes, seed):
    """ORBmatcher::SearchByBoW, both overloads (src/ORBmatcher.cc:155-263, :480-596); few nodes = long serial chains."""
    img, kps, desc = feats
    kf1, kf2, _, _, _ = synth.triangulation_case(kps, desc, seed, n_nodes=n_nodes)
    kf1["has_mp"] = 1 - kf1["has_mp"]; kf2["has_mp"] = 1 - kf2["has_mp"]
    e = O.search_by_bow(kf1, kf2, frame2, ratio, check_ori)
    g = matcher.search_by_bow(kf1, kf2, frame2, ratio, check_ori)
    assert g[0] == e[0] and e[0] > 50 and np.array_equal(g[1], e[1])


def test_bow_transform_feeds_search_by_bow(matcher, feats):
    """The chain the tracker runs: ComputeBoW on keyframe and frame, then SearchByBoW on the two FeatureVectors."""
    import bow_io
    img, kps, desc = feats
    t = bow_io.random_tree(10, 4, 3)
    voc = Y.Vocabulary.from_arrays(t["k"], t["L"], t["child_start"], t["child_list"], t["desc"], t["word_id"], t["weight"])
    ovoc = O.Vocabulary.from_arrays(t["k"], t["L"], t["child_start"], t["child_list"], t["desc"], t["word_id"], t["weight"])
    rng = np.random.default_rng(8)
    desc2 = synth.flip_bits(desc, rng.integers(0, 25, size=len(desc)), rng)
    kps2 = kps.copy(); kps2["angle"] = np.mod(kps2["angle"] + rng.normal(0, 5, len(kps)), 360).astype(np.float32)
    r1, r2 = Y.bow_transform(matcher, voc, [desc, desc2], 2)
    o1, o2 = ovoc.transform(desc, 2), ovoc.transform(desc2, 2)
    mk = lambda k, d, r: {"kps": k, "desc": d, "has_mp": np.ones(len(k), np.uint8), "node_ids": r["fv_node"],
                          "node_start": r["fv_start"], "node_feats": r["fv_feat"]}
    g = matcher.search_by_bow(mk(kps, desc, r1), mk(kps2, desc2, r2), True, 0.7, True)
    e = O.search_by_bow(mk(kps, desc, o1), mk(kps2, desc2, o2), True, 0.7, True)
    assert g[0] == e[0] and e[0] > 200 and np.array_equal(g[1], e[1])
    voc.close(); ovoc.close()


@pytest.mark.parametrize("w,h,nf,seed", [(640, 240, 800, 5), (1241, 376, 2000, 6)])
def test_stereo_vs_oracle(matcher, w, h, nf, seed):
    left, right = synth.stereo_pair(w, h, seed)
    exl, exr = Y.Extractor(nf), Y.Extractor(nf)
    kl, dl = exl.extract(left)
    kr, dr = exr.extract(right)
    p = O.params(nf)
    sf, isf = exl.tables()["scale"], exl.tables()["inv_scale"]
    bf, b = 386.1448, 386.1448 / 718.856                    # Examples/Stereo/KITTI00-02.yaml
    ur, depth, _ = O.stereo_match(kl, dl, kr, dr, O.pyramid(p, left), O.pyramid(p, right), sf, isf, bf, b)
    gur, gdepth = matcher.stereo_match(exl, exr, 0, kl, dl, kr, dr, bf, b)
    assert (ur >= 0).sum() > 100
    assert gur.tobytes() == ur.tobytes() and gdepth.tobytes() == depth.tobytes()
    exl.close(); exr.close()
