"""Runs oracle/_ref/ygz_ref_orb (the reference's own ORBextractor.cc built against oracle/minicv)."""
import json
import os
import subprocess
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "ygz_ref_orb")

LAST_STATS = None   # the JSON line of the last run_ref_dso call (frames, seconds: processing only)
KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])


def available():
    return os.path.exists(REF_BIN)


def run_ref(frames, nfeatures=1000, scale=1.2, levels=8, ini=20, mn=7, pyr=False, frame_api=False,
            repeat=1, threads=1):
    frames = np.ascontiguousarray(frames, np.uint8)
    if frames.ndim == 2:
        frames = frames[None]
    n, h, w = frames.shape
    with tempfile.TemporaryDirectory() as td:
        fin, fout = os.path.join(td, "in.u8"), os.path.join(td, "out.bin")
        frames.tofile(fin)
        cmd = [REF_BIN, "--in", fin, "--w", str(w), "--h", str(h), "--n", str(n), "--nfeatures", str(nfeatures),
               "--scale", repr(float(scale)), "--levels", str(levels), "--ini", str(ini), "--min", str(mn),
               "--out", fout, "--repeat", str(repeat), "--threads", str(threads)]
        if pyr:
            cmd.append("--pyr")
        if frame_api:
            cmd.append("--frame-api")
        res = subprocess.run(cmd, capture_output=True, text=True, check=True)
        stats = json.loads(res.stdout.strip().splitlines()[-1])
        buf = open(fout, "rb").read()
    outs, off = [], 0
    for _ in range(n):
        cnt = int(np.frombuffer(buf, "<i4", 1, off)[0]); off += 4
        kps = np.frombuffer(buf, KP_DTYPE, cnt, off).copy(); off += 28 * cnt
        desc = np.frombuffer(buf, np.uint8, 32 * cnt, off).reshape(cnt, 32).copy(); off += 32 * cnt
        pyrs = []
        if pyr:
            for _l in range(levels):
                lw, lh = np.frombuffer(buf, "<i4", 2, off); off += 8
                pyrs.append(np.frombuffer(buf, np.uint8, int(lw) * int(lh), off).reshape(int(lh), int(lw)).copy())
                off += int(lw) * int(lh)
        outs.append({"kps": kps, "desc": desc, "pyr": pyrs})
    return outs, stats


def run_ref_dso(frames, exist=None, nfeatures=1000, scale=1.2, levels=8, grid=-1, orbslam=False):
    """--dso mode: ONE extractor over the frames in order (mnGridSize carries over), DSO_KEYPOINT method on Frames that
    already hold exist[i] (KP_DTYPE arrays).  Returns per frame {"kps", "desc", "grid_after"}.
    orbslam=True: the same Frames through ORBSLAM_KEYPOINT, leftEye = true (src/ORBextractor.cc:1086-1099: descriptors of the
    existing keys in front of the new OctTree keys); grid_after is then absent."""
    frames = np.ascontiguousarray(frames, np.uint8)
    if frames.ndim == 2:
        frames = frames[None]
    n, h, w = frames.shape
    with tempfile.TemporaryDirectory() as td:
        fin, fout, fex = os.path.join(td, "in.u8"), os.path.join(td, "out.bin"), os.path.join(td, "exist.bin")
        frames.tofile(fin)
        cmd = [REF_BIN, "--in", fin, "--w", str(w), "--h", str(h), "--n", str(n), "--nfeatures", str(nfeatures),
               "--scale", repr(float(scale)), "--levels", str(levels), "--out", fout] + (["--frame-api"] if orbslam else ["--dso", "--grid", str(grid)])
        if exist is not None:
            with open(fex, "wb") as f:
                for e in exist:
                    np.array([len(e)], np.int32).tofile(f)
                    np.ascontiguousarray(e, KP_DTYPE).tofile(f)
            cmd += ["--exist", fex]
        res = subprocess.run(cmd, capture_output=True, text=True, check=True)
        global LAST_STATS
        LAST_STATS = json.loads(res.stdout.strip().splitlines()[-1])
        buf = open(fout, "rb").read()
    outs, off = [], 0
    for _ in range(n):
        cnt = int(np.frombuffer(buf, "<i4", 1, off)[0]); off += 4
        kps = np.frombuffer(buf, KP_DTYPE, cnt, off).copy(); off += 28 * cnt
        desc = np.frombuffer(buf, np.uint8, 32 * cnt, off).reshape(cnt, 32).copy(); off += 32 * cnt
        g = None
        if not orbslam:
            g = int(np.frombuffer(buf, "<i4", 1, off)[0]); off += 4
        outs.append({"kps": kps, "desc": desc, "grid_after": g})
    return outs
