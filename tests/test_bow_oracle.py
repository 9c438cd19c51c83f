"""Bag-of-words oracle (oracle/bow_oracle.cpp) pinned to the reference: against golden vectors produced by the
reference's own DBoW2 (tests/golden/bow_*.npz, tools/make_golden_bow.py) and, where oracle/_ref/ygz_ref_bow exists (the
authoring container and any box the snapshot travels to), against that binary run live on fresh seeded cases."""
import glob
import os
import tempfile

import numpy as np
import pytest

import bow_io
import oracle_lib as O

GOLDEN = sorted(glob.glob(os.path.join(O.ROOT, "tests", "golden", "bow_*.npz")))
KEYS = ["word", "node", "weight", "bow_word", "bow_value", "fv_node", "fv_start", "fv_feat"]


def _same(a, b, ctx):
    for k in KEYS:
        x, y = np.asarray(a[k]), np.asarray(b[k])
        assert x.shape == y.shape and x.tobytes() == y.astype(x.dtype).tobytes(), (ctx, k)   # doubles: bit for bit


def _load_golden_voc(g):
    with tempfile.NamedTemporaryFile(suffix=".bin", delete=False) as f:
        f.write(g["voc_bin"].tobytes())
        path = f.name
    try:
        return O.Vocabulary.load_binary(path)
    finally:
        os.unlink(path)


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[4:-4] for p in GOLDEN])
def test_restated_loader_and_transform_vs_reference_golden(path):
    g = np.load(path)
    voc = _load_golden_voc(g)
    a = voc.arrays()
    # the tree the reference holds after loadFromBinaryFile (incl. the duplicated last node of its read loop)
    assert voc.n_nodes == len(g["tree_parent"]) and voc.k == int(g["tree_k"]) and voc.L == int(g["tree_L"])
    for k in ("parent", "child_start", "child_list", "word_id"):
        assert np.array_equal(a[k], g["tree_" + k]), k
    assert np.array_equal(a["desc"][1:], g["tree_desc"][1:])
    assert a["weight"].tobytes() == g["tree_weight"].tobytes()
    assert a["child_list"][-1] == voc.n_nodes - 1 or voc.n_nodes - 1 in a["child_list"]   # the duplicate is somebody's child
    # same vocabulary through the flat-array constructor
    voc2 = O.Vocabulary.from_arrays(voc.k, voc.L, a["child_start"], a["child_list"], a["desc"], a["word_id"], a["weight"])
    for i in range(int(g["n_frames"])):
        want = {k: g["f%d_%s" % (i, k)] for k in KEYS}
        got = voc.transform(g["f%d_desc" % i], int(g["levelsup"]))
        _same(got, want, (path, i))
        _same(voc2.transform(g["f%d_desc" % i], int(g["levelsup"])), want, (path, i, "arrays"))
    voc.close(); voc2.close()


def test_golden_has_stop_words_and_multi_feature_words():
    """The fixtures exercise both BowVector branches: words hit by several features (value += w) and stopped words."""
    multi = stopped = 0
    for path in GOLDEN:
        g = np.load(path)
        for i in range(int(g["n_frames"])):
            w, wt = g["f%d_word" % i], g["f%d_weight" % i]
            stopped += int((wt <= 0).sum())
            live = w[wt > 0]
            multi += len(live) - len(np.unique(live))
    assert multi > 20 and stopped >= 0


@pytest.mark.skipif(not bow_io.have_ref(), reason="oracle/_ref/ygz_ref_bow not built (needs /root/reference)")
@pytest.mark.parametrize("k,L,levelsup,seed", [(10, 4, 2, 1), (5, 6, 4, 2), (8, 3, 3, 3)])
def test_restatement_vs_reference_live(k, L, levelsup, seed):
    rng = np.random.default_rng(seed)
    images = [rng.integers(0, 256, (60, 32), dtype=np.uint8) for _ in range(50)]
    pool = np.concatenate(images)
    frames = []
    for n in (1200, 333, 2):
        src = pool[rng.integers(0, len(pool), n)]
        frames.append(np.bitwise_xor(src, np.packbits(rng.uniform(size=(n, 256)) < 0.1, axis=1)))
    with tempfile.TemporaryDirectory() as td:
        vp = os.path.join(td, "voc.bin")
        bow_io.ref_create(k, L, 100 + seed, images, vp)
        want, _ = bow_io.ref_transform(vp, frames, levelsup)
        tree = bow_io.ref_dump(vp)
        voc = O.Vocabulary.load_binary(vp)
    a = voc.arrays()
    assert np.array_equal(a["child_list"], tree["child_list"]) and np.array_equal(a["word_id"], tree["word_id"])
    for i, f in enumerate(frames):
        _same(voc.transform(f, levelsup), want[i], (k, L, i))
    voc.close()


def test_random_tree_helper_is_a_valid_vocabulary():
    t = bow_io.random_tree(6, 4, 5, p_stop=0.15)
    voc = O.Vocabulary.from_arrays(t["k"], t["L"], t["child_start"], t["child_list"], t["desc"], t["word_id"], t["weight"])
    f = bow_io.features_near_nodes(t, 500, 6)
    r = voc.transform(f, 2)
    assert len(r["bow_word"]) > 50 and abs(r["bow_value"].sum() - 1.0) < 1e-12      # L1-normalised
    assert np.all(np.diff(r["bow_word"]) > 0) and np.all(np.diff(r["fv_node"]) > 0)
    assert sorted(r["fv_feat"].tolist()) == np.nonzero(r["weight"] > 0)[0].tolist()
    voc.close()
