"""First-contact diagnostic run on the GPU box: stage-by-stage comparison with verbose output."""
import os
import sys
import traceback

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import golden_cases  # noqa: E402
import oracle_lib as O  # noqa: E402
from ygz import Y  # noqa: E402


def main():
    g = golden_cases.load(sys.argv[1] if len(sys.argv) > 1 else "small200")
    img = g["img"]; h, w = img.shape
    p = O.params(g["nfeatures"], g["scale"], g["levels"], g["ini"], g["min"])
    ex = Y.Extractor(g["nfeatures"], g["scale"], g["levels"], g["ini"], g["min"])
    kps, desc = ex.extract(img, Y.KEEP_STAGES)
    opyr = O.pyramid(p, img)
    pyr = ex.pyramid(0, w, h)
    for l in range(g["levels"]):
        print("level", l, pyr[l].shape, "pyr mismatches", int((pyr[l] != opyr[l]).sum()))
        xs, ys, rs = ex.candidates(0, l, w, h)
        oxs, oys, ors = O.fast_candidates(opyr[l], g["ini"], g["min"])
        same = len(xs) == len(oxs) and np.array_equal(xs, oxs) and np.array_equal(ys, oys) and np.array_equal(rs, ors)
        print("   cand", len(xs), len(oxs), "same" if same else "DIFF")
        bl = ex.blurred_level(0, l, w, h)
        print("   blur mismatches", int((bl != O.blur7(opyr[l])).sum()))
    print("n", len(kps), len(g["kps"]))
    n = min(len(kps), len(g["kps"]))
    for fld in ("x", "y", "size", "angle", "response", "octave", "class_id"):
        print("  ", fld, "mismatch", int((kps[fld][:n] != g["kps"][fld][:n]).sum()))
    print("   desc rows mismatch", int((desc[:n] != g["desc"][:n]).any(axis=1).sum()))
    print("per-level counts", np.bincount(kps["octave"], minlength=g["levels"]).tolist(),
          np.bincount(g["kps"]["octave"], minlength=g["levels"]).tolist())


if __name__ == "__main__":
    try:
        main()
    except Exception:
        traceback.print_exc()
        sys.exit(1)
