"""Golden vectors of the bag-of-words assignment, made by the reference's OWN DBoW2 (oracle/_ref/ygz_ref_bow, built by
oracle/Makefile from /root/reference/Thirdparty/DBoW2): a vocabulary trained by TemplatedVocabulary::create, saved by
saveToBinaryFile, and the outputs of loadFromBinaryFile + transform on seeded descriptors.
    python tools/make_golden_bow.py      -> tests/golden/bow_*.npz"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bow_io  # noqa: E402

CASES = [  # name, k, L, create seed, training images x descriptors each, levelsup, frames (descriptor counts)
    ("k6L4", 6, 4, 7, 40, 30, 2, (300, 257, 1, 0)),
    ("k10L3", 10, 3, 11, 60, 50, 4, (400, 64)),
    ("k4L5", 4, 5, 3, 25, 40, 3, (350,)),
]


def main():
    assert bow_io.have_ref(), "build oracle/_ref first (make -C oracle ref)"
    for name, k, L, seed, n_img, per_img, levelsup, counts in CASES:
        rng = np.random.default_rng(1000 + seed)
        # uniformly random training descriptors, few per leaf: the tree comes out irregular (branches that stop early).
        # (Tightly clustered data makes the reference's own k-means crash on an empty cluster, so it is avoided.)
        images = [rng.integers(0, 256, (per_img, 32), dtype=np.uint8) for _ in range(n_img)]
        with tempfile.TemporaryDirectory() as td:
            vp = os.path.join(td, "voc.bin")
            print(bow_io.ref_create(k, L, seed, images, vp).strip())
            voc_bytes = np.frombuffer(open(vp, "rb").read(), np.uint8)
            pool = np.concatenate(images)
            frames = []
            for n in counts:
                src = pool[rng.integers(0, len(pool), n)] if n else np.zeros((0, 32), np.uint8)
                flips = (rng.uniform(size=(n, 256)) < 0.08)
                frames.append(np.bitwise_xor(src, np.packbits(flips, axis=1)) if n else src)
            res, _ = bow_io.ref_transform(vp, frames, levelsup)
            tree = bow_io.ref_dump(vp)
        out = {"voc_bin": voc_bytes, "levelsup": np.int32(levelsup), "n_frames": np.int32(len(frames))}
        for key, val in tree.items():
            out["tree_" + key] = np.asarray(val)
        for i, (f, r) in enumerate(zip(frames, res)):
            out["f%d_desc" % i] = f
            for key, val in r.items():
                out["f%d_%s" % (i, key)] = val
        path = os.path.join(ROOT, "tests", "golden", "bow_%s.npz" % name)
        np.savez_compressed(path, **out)
        print(path, os.path.getsize(path), "bytes; nodes", len(tree["parent"]), "bow sizes", [len(r["bow_word"]) for r in res])


if __name__ == "__main__":
    main()
