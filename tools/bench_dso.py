"""DSO_KEYPOINT extraction through the C ABI with host buffers: B independent 752x480 frames per call,
150 existing keypoints each, next to the reference's own code (oracle/_ref/ygz_ref_orb --dso, single thread — the mode is
sequential per camera stream in the reference).  Results of the first frames are checked against the reference.
One JSON line."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb-ygz-slam_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import refcli  # noqa: E402
import synth  # noqa: E402
import ygzorb_ctypes as Y  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
    w, h, nf = 752, 480, 1000
    distinct = 8
    base = np.stack([synth.textured_frame(w, h, 300 + k) for k in range(distinct)])
    frames = base[np.arange(B) % distinct]
    rng = np.random.default_rng(1)
    exist = []
    for k in range(distinct):
        e = np.zeros(150, Y.KP_DTYPE)
        e["x"] = rng.uniform(40, w - 40, 150).astype(np.float32); e["y"] = rng.uniform(40, h - 40, 150).astype(np.float32)
        e["octave"] = rng.integers(0, 4, 150); e["size"] = 7.0; e["angle"] = -1.0; e["class_id"] = -1
        exist.append(e)
    ex_all = [exist[k % distinct] for k in range(B)]
    grids = [23] * B                       # the steady-state grid of this texture (avoids the warm-up iterations)
    ex = Y.Extractor(nf, max_batch=B)
    CAP = 4096
    got, ga = ex.extract_dso(frames, ex_all, grids, cap=CAP)
    ok = None
    if refcli.available():
        ok = True
        for k in range(distinct):
            r = refcli.run_ref_dso(frames[k], [exist[k]], nfeatures=nf, grid=23)[0]
            ok &= got[k][0].tobytes() == r["kps"].tobytes() and np.array_equal(got[k][1], r["desc"]) and ga[k] == r["grid_after"]
    reps = 5
    t0 = time.perf_counter()
    for _ in range(reps):
        ex.extract_dso(frames, ex_all, grids, cap=CAP)
    gpu = B * reps / (time.perf_counter() - t0)
    cpu = None
    if refcli.available():
        sec = 0.0
        for k in range(distinct):
            refcli.run_ref_dso(frames[k], [exist[k]], nfeatures=nf, grid=23)
            sec += refcli.LAST_STATS["seconds"]          # processing only (no process start, no file I/O)
        cpu = distinct / sec
    print(json.dumps({"bench": "DSO_KEYPOINT extraction, host buffers (wrapper call incl. packing)", "frames_per_call": B,
                      "gpu_frames_per_s": gpu, "cpu_reference_frames_per_s_1thread": cpu, "matches_reference": ok,
                      "keypoints_per_frame": float(np.mean([len(g[0]) for g in got]))}))
    Y.close_all()   # orderly shutdown: handles first, then a normal interpreter exit (exit hooks run)


if __name__ == "__main__":
    main()
