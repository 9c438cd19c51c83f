"""Per-stage device times of single-frame extraction (the live-tracking shape of configs 0-2): python tools/latency_stages.py [w h nfeat]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb-ygz-slam_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import synth  # noqa: E402
import ygzorb_ctypes as Y  # noqa: E402

w, h, nf = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (752, 480, 1000)
img = synth.textured_frame(w, h, 3)
ex = Y.Extractor(nf)
for _ in range(20):
    ex.extract(img)
t0 = time.perf_counter()
for _ in range(200):
    ex.extract(img)
wall = (time.perf_counter() - t0) / 200
ex.set_profiling(True)
acc = np.zeros(8)
for _ in range(50):
    ex.extract(img)
    acc += ex.stage_ms()
names = ["upload", "pyramid", "fast", "quadtree", "blur", "orient_desc", "download"]
print({"frame": [w, h, nf], "wall_ms_per_call": wall * 1e3, "stage_us": {n: float(acc[i] / 50 * 1e3) for i, n in enumerate(names)}})
