"""File formats of oracle/_ref/ygz_ref_bow (the reference's own DBoW2, see oracle/bow_ref_cli.cpp) and seeded
synthetic vocabularies / feature sets for the bag-of-words tests."""
import os
import subprocess
import tempfile

import numpy as np

import oracle_lib as O


def have_ref():
    return os.path.exists(O.REF_BOW)


def write_training(path, images):
    with open(path, "wb") as f:
        np.array([len(images)], np.int32).tofile(f)
        for d in images:
            np.array([len(d)], np.int32).tofile(f)
            np.ascontiguousarray(d, np.uint8).tofile(f)


def write_features(path, frames):
    write_training(path, frames)


def ref_create(k, L, seed, images, voc_path):
    """TemplatedVocabulary::create of the reference on `images` (list of (n,32) uint8) -> voc_path (binary format)."""
    with tempfile.TemporaryDirectory() as td:
        tp = os.path.join(td, "train.bin")
        write_training(tp, images)
        r = subprocess.run([O.REF_BOW, "create", str(k), str(L), str(seed), tp, voc_path], capture_output=True, text=True)
        assert r.returncode == 0, r.stdout + r.stderr
        return r.stdout


def parse_transform_out(buf, n_frames):
    out, off = [], 0

    def take(dt, n):
        nonlocal off
        a = np.frombuffer(buf, dt, n, off)
        off += a.nbytes
        return a
    rec = np.dtype([("word", "<i4"), ("node", "<i4"), ("weight", "<f8")])
    brec = np.dtype([("word", "<i4"), ("value", "<f8")])
    for _ in range(n_frames):
        n = int(take("<i4", 1)[0])
        r = take(rec, n)
        nb = int(take("<i4", 1)[0])
        b = take(brec, nb)
        nn = int(take("<i4", 1)[0])
        nodes, start, feats = [], [0], []
        for _ in range(nn):
            h = take("<i4", 2)
            nodes.append(int(h[0]))
            feats.append(take("<i4", int(h[1])))
            start.append(start[-1] + int(h[1]))
        out.append({"word": r["word"].copy(), "node": r["node"].copy(), "weight": r["weight"].copy(),
                    "bow_word": b["word"].copy(), "bow_value": b["value"].copy(),
                    "fv_node": np.array(nodes, np.int32), "fv_start": np.array(start, np.int32),
                    "fv_feat": np.concatenate(feats).astype(np.int32) if feats else np.zeros(0, np.int32)})
    return out


def ref_transform(voc_path, frames, levelsup, threads=1, repeat=1):
    """The reference's loadFromBinaryFile + transform on every frame.  Returns (list of result dicts, stdout)."""
    with tempfile.TemporaryDirectory() as td:
        fp, op = os.path.join(td, "feats.bin"), os.path.join(td, "out.bin")
        write_features(fp, frames)
        r = subprocess.run([O.REF_BOW, "transform", voc_path, str(levelsup), fp, op, str(threads), str(repeat)],
                           capture_output=True, text=True)
        assert r.returncode == 0, r.stdout + r.stderr
        return parse_transform_out(open(op, "rb").read(), len(frames)), r.stdout


def ref_dump(voc_path):
    """The tree the reference holds after loadFromBinaryFile, as flat arrays."""
    with tempfile.TemporaryDirectory() as td:
        dp = os.path.join(td, "flat.bin")
        r = subprocess.run([O.REF_BOW, "dump", voc_path, dp], capture_output=True, text=True)
        assert r.returncode == 0, r.stdout + r.stderr
        buf = open(dp, "rb").read()
    n, k, L = [int(x) for x in np.frombuffer(buf, "<i4", 3)]
    rec = np.dtype([("parent", "<i4"), ("nch", "<i4"), ("word", "<i4"), ("weight", "<f8"), ("desc", "u1", 32)])
    nodes = np.frombuffer(buf, rec, n, 12)
    ch = np.frombuffer(buf, "<i4", int(nodes["nch"].sum()), 12 + n * rec.itemsize)
    start = np.concatenate([[0], np.cumsum(nodes["nch"])]).astype(np.int32)
    return {"k": k, "L": L, "parent": nodes["parent"].copy(), "child_start": start, "child_list": ch.copy(),
            "desc": nodes["desc"].copy(), "word_id": nodes["word"].copy(), "weight": nodes["weight"].copy()}


def random_tree(k, L, seed, p_stop=0.0, p_zero=0.05):
    """Seeded synthetic vocabulary as flat arrays: a k-ary tree of depth L whose inner nodes stop early with probability
    p_stop (irregular depth, fewer than k children), random node descriptors, float-valued idf-like weights with a
    fraction p_zero of stop words (weight 0).  transform() does not care how a tree was trained."""
    rng = np.random.default_rng(seed)
    children = [[]]
    depth = [0]
    frontier = [0]
    while frontier:
        nxt = []
        for nid in frontier:
            if depth[nid] >= L or (nid != 0 and rng.uniform() < p_stop):
                continue
            nk = k if rng.uniform() >= p_stop else int(rng.integers(1, k + 1))
            for _ in range(nk):
                children.append([]); depth.append(depth[nid] + 1)
                children[nid].append(len(children) - 1)
                nxt.append(len(children) - 1)
        frontier = nxt
    n = len(children)
    desc = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    desc[0] = 0
    word = np.full(n, -1, np.int32)
    leaves = [i for i in range(n) if not children[i]]
    word[leaves] = np.arange(len(leaves), dtype=np.int32)
    weight = np.zeros(n, np.float64)
    w = rng.uniform(0.1, 9.0, len(leaves)).astype(np.float32).astype(np.float64)
    w[rng.uniform(size=len(leaves)) < p_zero] = 0.0
    weight[leaves] = w
    start = np.zeros(n + 1, np.int32)
    for i in range(n):
        start[i + 1] = start[i] + len(children[i])
    clist = np.array([c for ch in children for c in ch], np.int32)
    return {"k": k, "L": L, "child_start": start, "child_list": clist, "desc": desc, "word_id": word, "weight": weight}


def features_near_nodes(tree, n, seed, flips=30):
    """n descriptors: random leaves' descriptors with up to `flips` random bit flips (so that descents are contested)."""
    import synth
    rng = np.random.default_rng(seed)
    leaves = np.nonzero(tree["word_id"] >= 0)[0]
    src = rng.choice(leaves, size=n)
    return synth.flip_bits(tree["desc"][src], rng.integers(0, flips + 1, size=n), rng)
