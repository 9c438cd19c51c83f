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


def full_tree(k, L, seed, p_zero=0.02):
    """Full k-ary tree of depth L in BFS NodeId order (the shape of ORBvoc: k=10, L=6 -> 1,111,111 nodes), vectorised."""
    rng = np.random.default_rng(seed)
    sizes = [k ** l for l in range(L + 1)]
    n = sum(sizes)
    n_inner = n - sizes[-1]
    child_start = np.zeros(n + 1, np.int64)
    child_start[1:n_inner + 1] = np.arange(1, n_inner + 1) * k
    child_start[n_inner + 1:] = n_inner * k
    child_list = np.arange(1, n, dtype=np.int32)
    desc = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    desc[0] = 0
    word = np.full(n, -1, np.int32)
    word[n_inner:] = np.arange(sizes[-1], dtype=np.int32)
    weight = np.zeros(n, np.float64)
    w = rng.uniform(0.1, 9.0, sizes[-1]).astype(np.float32).astype(np.float64)
    w[rng.uniform(size=sizes[-1]) < p_zero] = 0.0
    weight[n_inner:] = w
    parent = np.zeros(n, np.int32)
    parent[1:] = (np.arange(1, n) - 1) // k
    return {"k": k, "L": L, "child_start": child_start.astype(np.int32), "child_list": child_list, "desc": desc, "word_id": word,
            "weight": weight, "parent": parent}


def write_binary_voc(path, tree):
    """TemplatedVocabulary::saveToBinaryFile layout (TemplatedVocabulary.h:1528-1550): header nb_nodes, size_node, k, L,
    scoring, weighting; per node 1..n-1: int32 parent, 32 B descriptor, float weight, bool isLeaf.  Needs tree['parent']."""
    n = len(tree["word_id"])
    rec = np.zeros(n - 1, np.dtype([("parent", "<i4"), ("desc", "u1", 32), ("weight", "<f4"), ("leaf", "u1")]))
    rec["parent"] = tree["parent"][1:]
    rec["desc"] = tree["desc"][1:]
    rec["weight"] = tree["weight"][1:].astype(np.float32)
    rec["leaf"] = (tree["child_start"][1:n] == tree["child_start"][2:n + 1]).astype(np.uint8)
    with open(path, "wb") as f:
        np.array([n, rec.dtype.itemsize], np.uint32).tofile(f)
        np.array([tree["k"], tree["L"], 0, 0], np.int32).tofile(f)      # L1_NORM = 0, TF_IDF = 0 (BowVector.h:30-47)
        rec.tofile(f)
