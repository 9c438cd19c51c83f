"""Loader for tests/golden/extract_*.npz (outputs of the reference's own ORBextractor.cc, made by
tools/make_golden.py).  The frames are regenerated from their seeds and checked against a CRC."""
import glob
import os
import zlib

import numpy as np

import synth

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def names():
    return sorted(os.path.basename(p)[len("extract_"):-4] for p in glob.glob(os.path.join(GOLDEN_DIR, "extract_*.npz")))


def load(name):
    z = np.load(os.path.join(GOLDEN_DIR, "extract_%s.npz" % name))
    w, h, seed, nf, lv, ini, mn = [int(v) for v in z["meta"]]
    kind = str(z["kind"])
    if kind == "textured":
        img = synth.textured_frame(w, h, seed)
    elif kind == "mixed":
        img = synth.mixed_frame(w, h, seed)
    else:
        img = synth.textured_frame(w, h, seed, contrast=0.1)
    assert zlib.crc32(img.tobytes()) == int(z["frame_crc32"]), "synthetic generator drifted from the fixture"
    return {"img": img, "nfeatures": nf, "scale": float(z["scale"]), "levels": lv, "ini": ini, "min": mn,
            "kps": z["kps"], "desc": z["desc"], "pyr_crc32": z["pyr_crc32"], "pyr_sizes": z["pyr_sizes"]}
