"""Generates tests/golden/dso_*.npz with the reference's OWN code (oracle/_ref/ygz_ref_orb --dso = src/ORBextractor.cc +
Thirdparty/fast compiled unmodified): a 3-frame stream per case, with existing keypoints.  The frames are regenerated from their seeds by
tests/synth.py (their CRC32s are stored).  usage: python tools/make_golden_dso.py"""
import os
import sys
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import refcli  # noqa: E402
import synth  # noqa: E402

CASES = [("dso_euroc752", 752, 480, 1000, "textured", 0), ("dso_mixed640", 640, 480, 800, "mixed", 5)]


def main():
    for name, w, h, nf, kind, seed in CASES:
        mk = synth.textured_frame if kind == "textured" else synth.mixed_frame
        frames = np.stack([mk(w, h, seed + k) for k in range(3)])
        first = refcli.run_ref_dso(frames, nfeatures=nf)
        rng = np.random.default_rng(seed)
        exist = []
        for k in range(3):
            n = 120
            e = np.zeros(n, refcli.KP_DTYPE)
            e["x"] = rng.uniform(40, w - 40, n).astype(np.float32); e["y"] = rng.uniform(40, h - 40, n).astype(np.float32)
            e["octave"] = rng.integers(0, 4, n); e["size"] = 7.0; e["angle"] = -1.0; e["class_id"] = -1
            pick = rng.choice(len(first[k]["kps"]), 60, replace=False)
            e["x"][:60] = first[k]["kps"]["x"][pick]; e["y"][:60] = first[k]["kps"]["y"][pick]; e["octave"][:60] = 0
            exist.append(e)
        outs = refcli.run_ref_dso(frames, exist, nfeatures=nf)
        d = {"w": np.int32(w), "h": np.int32(h), "seed": np.int32(seed), "textured": np.int32(kind == "textured"),
             "frame_crc32": np.array([zlib.crc32(f.tobytes()) for f in frames], np.uint32), "nfeatures": np.int32(nf), "grid_after": np.array([o["grid_after"] for o in outs], np.int32)}
        for k in range(3):
            d["exist_%d" % k] = exist[k]; d["kps_%d" % k] = outs[k]["kps"]; d["desc_%d" % k] = outs[k]["desc"]
        np.savez_compressed(os.path.join(ROOT, "tests", "golden", name + ".npz"), **d)
        print(name, [len(o["kps"]) for o in outs], d["grid_after"])


if __name__ == "__main__":
    main()
