"""Generates tests/golden/extract_*.npz by running the reference's OWN ORBextractor.cc
(oracle/_ref/ygz_ref_orb, built by oracle/Makefile from /root/reference) on seeded synthetic frames.
Needs the authoring container (reference tree present).  usage: python tools/make_golden.py
"""
import os
import sys
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import refcli  # noqa: E402
import synth  # noqa: E402

CASES = [
    # name, w, h, seed, kind, nfeatures, scale, levels, ini, min
    ("tum640", 640, 480, 7, "textured", 1000, 1.2, 8, 20, 7),
    ("euroc752", 752, 480, 0, "textured", 1000, 1.2, 8, 20, 7),
    ("euroc752_yaml", 752, 480, 0, "textured", 1000, 2.0, 4, 20, 7),
    ("kitti1241", 1241, 376, 3, "textured", 2000, 1.2, 8, 20, 7),
    ("hd1280", 1280, 720, 5, "textured", 2000, 1.2, 8, 20, 7),
    ("mixed752", 752, 480, 11, "mixed", 1000, 1.2, 8, 20, 7),
    ("lowc640", 640, 480, 21, "lowcontrast", 1000, 1.2, 8, 20, 7),
    ("small200", 200, 150, 4, "textured", 300, 1.2, 8, 20, 7),
]


def make_frame(kind, w, h, seed):
    if kind == "textured":
        return synth.textured_frame(w, h, seed)
    if kind == "mixed":
        return synth.mixed_frame(w, h, seed)
    if kind == "lowcontrast":
        return synth.textured_frame(w, h, seed, contrast=0.1)
    raise ValueError(kind)


def main():
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    for name, w, h, seed, kind, nf, sc, lv, ini, mn in CASES:
        img = make_frame(kind, w, h, seed)
        outs, _ = refcli.run_ref(img, nfeatures=nf, scale=sc, levels=lv, ini=ini, mn=mn, pyr=True)
        o = outs[0]
        crcs = np.array([zlib.crc32(p.tobytes()) for p in o["pyr"]], np.uint32)
        sizes = np.array([[p.shape[1], p.shape[0]] for p in o["pyr"]], np.int32)
        np.savez_compressed(os.path.join(out_dir, "extract_%s.npz" % name),
                            kps=o["kps"], desc=o["desc"], pyr_crc32=crcs, pyr_sizes=sizes,
                            frame_crc32=np.uint32(zlib.crc32(img.tobytes())),
                            meta=np.array([w, h, seed, nf, lv, ini, mn], np.int32), scale=np.float32(sc),
                            kind=np.array(kind))
        print(name, len(o["kps"]), np.bincount(o["kps"]["octave"], minlength=lv).tolist())


if __name__ == "__main__":
    main()
