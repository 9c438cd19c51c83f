"""The ORBmatcher searches of SURVEY.md 8(f) f2 through the C ABI with HOST buffers (copies inside the timed call), batched
over PAIRS frame pairs of ~1000 keypoints each, next to the oracle port on the host cores (threads over pairs).
Every GPU result is checked against the oracle on the first pairs before timing.  One JSON line."""
import argparse
import ctypes as C
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb-ygz-slam_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as O  # noqa: E402
import synth  # noqa: E402
import ygzorb_ctypes as Y  # noqa: E402


def timed(fn, reps):
    fn()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    return (time.perf_counter() - t0) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pairs", type=int, default=256)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--cpu-pairs", type=int, default=32)
    args = ap.parse_args()
    B = args.pairs
    w, h = 752, 480
    img = synth.textured_frame(w, h, 5)
    kps, desc = O.extract(O.params(1000), img)
    n = len(kps)
    sf, _, sigma2 = synth.scale_tables()
    inv_sigma2 = (np.float32(1.0) / sigma2).astype(np.float32)
    bounds = (0.0, 0.0, float(w), float(h))
    mt = Y.Matcher(0)
    cores = os.cpu_count() or 1
    pool = ThreadPoolExecutor(cores)
    out = {}
    distinct = 8     # distinct synthetic cases, cycled to fill the batch

    def cpu_rate(fn):
        k = args.cpu_pairs
        t0 = time.perf_counter()
        list(pool.map(fn, range(k)))
        return k / (time.perf_counter() - t0)

    names = {0: "SearchByProjection(Frame,KeyFrame) reloc", 1: "SearchByProjection(KeyFrame,Scw) sim3", 2: "Fuse(KeyFrame,points)",
             3: "Fuse(KeyFrame,Scw,points)"}
    for variant in (0, 1, 2, 3):
        cases = [synth.keyframe_points_case(kps, desc, 1000, 100 + 10 * variant + k, variant) for k in range(distinct)]
        frames = [(kps, desc, cases[k % distinct][1], cases[k % distinct][0]["u_right"] if variant == 2 else None) for k in range(B)]
        pts = [cases[k % distinct][0] for k in range(B)]
        th, od, ori = (10.0, 100, True) if variant == 0 else ((10.0, 50, False) if variant == 1 else (3.0, 50, False))
        keep = []
        F, L, n_kp, n_pt, ks, ps = mt._keyframe_views(frames, pts, bounds, sf, keep)
        kp_to_pt = np.zeros((B, ks), np.int32); best = np.zeros((B, ps), np.int32); nm = np.zeros(B, np.int32)
        p = lambda a: C.c_void_p(a.ctypes.data)
        if variant == 0:
            call = lambda: Y.check(Y.lib().ygzorb_search_by_projection_reloc(mt.h, C.byref(F), C.byref(L), th, od, 1, 0, p(kp_to_pt), p(best), p(nm)))
        elif variant == 1:
            call = lambda: Y.check(Y.lib().ygzorb_search_by_projection_sim3(mt.h, C.byref(F), C.byref(L), int(th), 0, p(kp_to_pt), p(best), p(nm)))
        else:
            call = lambda: Y.check(Y.lib().ygzorb_fuse(mt.h, C.byref(F), C.byref(L), th, int(variant == 3), p(inv_sigma2), 0, p(best), p(nm)))
        sec = timed(call, args.reps)

        def one(k, variant=variant, th=th, od=od, ori=ori):
            c, cl = cases[k % distinct]
            return O.search_by_projection_keyframe(kps, desc, cl if variant <= 1 else None, bounds, sf, c, th, variant, od, ori,
                                                   c["u_right"] if variant == 2 else None, inv_sigma2)
        for k in range(distinct):
            e = one(k)
            assert int(nm[k]) == e[0] and np.array_equal(best[k, :n_pt[k]], e[2]), names[variant]
        out[names[variant]] = {"gpu_pairs_per_s": B / sec, "gpu_ms_per_batch": sec * 1e3, "cpu_pairs_per_s": cpu_rate(one),
                               "matches_per_pair": float(nm.mean())}

    # SearchForInitialization
    cases = [synth.initialization_case(kps, desc, 200 + k) for k in range(distinct)]
    f1 = [(cases[k % distinct][0], cases[k % distinct][1]) for k in range(B)]
    f2 = [(cases[k % distinct][2], cases[k % distinct][3]) for k in range(B)]
    prevs = [cases[k % distinct][4] for k in range(B)]
    t0 = time.perf_counter()
    got = mt.search_for_initialization(f1, f2, prevs, bounds, 100, 0.9, True)
    wrap = time.perf_counter() - t0
    t0 = time.perf_counter()
    got = mt.search_for_initialization(f1, f2, prevs, bounds, 100, 0.9, True)
    wrap = time.perf_counter() - t0

    def one_init(k):
        c = cases[k % distinct]
        return O.search_for_initialization(c[0], c[1], c[2], c[3], bounds, c[4], 100, 0.9, True)
    for k in range(distinct):
        e = one_init(k)
        assert got[k][0] == e[0] and np.array_equal(got[k][1], e[1])
    out["SearchForInitialization"] = {"gpu_pairs_per_s": B / wrap, "gpu_ms_per_batch": wrap * 1e3, "cpu_pairs_per_s": cpu_rate(one_init),
                                      "matches_per_pair": float(np.mean([g[0] for g in got])),
                                      "note": "GPU time includes the Python packing of the batch (wrapper call)"}
    print(json.dumps({"bench": "f2 matcher searches, host buffers through the C ABI", "pairs": B, "keypoints": n, "cpu_threads": cores,
                      "results": out}))
    Y.close_all()   # orderly shutdown: handles first, then a normal interpreter exit (exit hooks run)


if __name__ == "__main__":
    main()
