#!/usr/bin/env python3
"""Runs the REFERENCE'S OWN matcher code (oracle/_ref/libygz_ref_match.so = /root/reference/src/ORBmatcher.cc + src/Align.cc +
the hot-path bodies of Frame / KeyFrame / MapPoint, compiled unmodified; oracle/Makefile target refmatch) on every case of
tests/match_cases.py and commits its outputs as tests/golden/match_golden.npz.  Run in the authoring container only (needs
/root/reference at build time).  Entries: "<case key>#o<i>" = output i; "<case key>#d_<name>" = an input only the reference
can derive (decomposed Sim3 pose, sR12 / sR21 / t21, TCR); "<case key>#in" = sha1 of the case's seeded inputs, so that a
generator drifting between machines is told apart from a parity failure."""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import match_cases as MC  # noqa: E402
import ref_match_lib as R  # noqa: E402


def digest(obj, h=None):
    top = h is None
    h = h or hashlib.sha1()
    if isinstance(obj, dict):
        for k in sorted(obj):
            h.update(k.encode()); digest(obj[k], h)
    elif isinstance(obj, (list, tuple)):
        for v in obj:
            digest(v, h)
    elif obj is None:
        h.update(b"None")
    else:
        a = np.ascontiguousarray(obj)
        h.update(str(a.dtype).encode()); h.update(a.tobytes())
    return h.hexdigest() if top else None


def main():
    if not R.available():
        raise SystemExit("oracle/_ref/libygz_ref_match.so is missing: run `make -C oracle refmatch` where /root/reference exists")
    out = {}
    for case in MC.CASES:
        d = MC.build_inputs(case)
        outs, der = MC.run_ref(d)
        key = case[1]
        out[key + "#in"] = np.frombuffer(bytes.fromhex(digest(d)), np.uint8)
        for i, a in enumerate(outs):
            out["%s#o%d" % (key, i)] = np.asarray(a)
        for name, a in der.items():
            out["%s#d_%s" % (key, name)] = np.asarray(a)
    path = os.path.join(ROOT, "tests", "golden", "match_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote %s: %d cases, %d arrays, %.1f KB" % (path, len(MC.CASES), len(out), os.path.getsize(path) / 1e3))


if __name__ == "__main__":
    main()
