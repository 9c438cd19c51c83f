#!/usr/bin/env python3
"""Single-frame (live tracking) latency of BASELINE.json configs[0] and configs[2] THROUGH the drop-in C++ class
(orb-ygz-slam_b200/host/host_latency: ygz::ORBextractor::operator() with a pageable image in, keypoints, descriptors and
mvImagePyramid out; stereo: Left / Right extractors on two host threads + ComputeStereoMatches), with the reference's own CPU
code timed beside it on the same box (oracle/_ref/ygz_ref_orb = src/ORBextractor.cc, one thread per image as the reference
runs it; stereo adds Frame::ComputeStereoMatches from oracle/_ref/libygz_ref_match.so).  One JSON line per config on stdout."""
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as O  # noqa: E402
import ref_match_lib as R  # noqa: E402
import refcli  # noqa: E402
import synth  # noqa: E402

LAT = os.path.join(ROOT, "orb-ygz-slam_b200", "host", "host_latency")


def run_host(frames, w, h, nf, reps, stereo):
    with tempfile.TemporaryDirectory() as td:
        fin = os.path.join(td, "in.u8")
        np.ascontiguousarray(frames, np.uint8).tofile(fin)
        cmd = [LAT, fin, str(w), str(h), str(nf), str(reps)] + (["stereo"] if stereo else [])
        out = subprocess.run(cmd, capture_output=True, text=True, check=True).stdout
    return json.loads(out.strip().splitlines()[-1])


def main():
    res = []
    # configs[0]: one 640x480 frame, 1000 features
    img = synth.textured_frame(640, 480, 7)
    g = run_host(img, 640, 480, 1000, 300, False)
    _, st = refcli.run_ref(img[None], nfeatures=1000, repeat=40, threads=1)
    res.append({"config": "configs[0]: ORBextractor on one 640x480 frame, 1000 features, scale 1.2, 8 levels, FAST 20/7", "gpu": g,
                "cpu_baseline": {"kind": "reference", "cores": 1, "latency_ms": 1e3 / st["frames_per_s"],
                                 "sample": "oracle/_ref/ygz_ref_orb (src/ORBextractor.cc), 1 thread, 40 repeats"},
                "speedup": (1e3 / st["frames_per_s"]) / (g["latency_us_median"] * 1e-3)})
    # configs[2]: 1241x376 stereo pair, 2000 features per image + ComputeStereoMatches
    left, right = synth.stereo_pair(1241, 376, 6)
    g = run_host(np.stack([left, right]), 1241, 376, 2000, 200, True)
    _, st = refcli.run_ref(np.stack([left, right]), nfeatures=2000, repeat=20, threads=2)      # Left / Right threads (src/Frame.cc:728-731)
    pair_ms = 2e3 / st["frames_per_s"]
    stereo_ms, kind = None, "reference"
    p = O.params(2000)
    kl, dl = O.extract(p, left); kr, dr = O.extract(p, right)
    sf, isf, _ = synth.scale_tables()
    pl, pr = O.pyramid(p, left), O.pyramid(p, right)
    fn = R.stereo_match if R.available() else (lambda *a: O.stereo_match(*a)[:2])
    kind2 = "reference" if R.available() else "oracle-port"
    t0 = time.perf_counter()
    for _ in range(20):
        fn(kl, dl, kr, dr, pl, pr, sf, isf, 386.1448, 386.1448 / 718.856)
    stereo_ms = (time.perf_counter() - t0) / 20 * 1e3
    res.append({"config": "configs[2]: 1241x376 stereo pair, 2000 features per image + Frame::ComputeStereoMatches", "gpu": g,
                "cpu_baseline": {"kind": kind, "kind_stereo_match": kind2, "cores": 2, "latency_ms": pair_ms + stereo_ms,
                                 "extract_pair_ms": pair_ms, "stereo_match_ms": stereo_ms,
                                 "sample": "ygz_ref_orb on the two images with 2 threads (20 repeats) + ComputeStereoMatches x 20 "
                                           "(the call includes building the Frame from flat arrays)"},
                "speedup": (pair_ms + stereo_ms) / (g["latency_us_median"] * 1e-3)})
    for r in res:
        print(json.dumps(r), flush=True)


if __name__ == "__main__":
    main()
