"""GPU parity of the bag-of-words assignment (ygzorb_vocabulary_* / ygzorb_bow_transform) against the oracle, whose
restatement is pinned to the reference's own DBoW2 (tests/test_bow_oracle.py), and directly against the golden vectors
that DBoW2 produced: words, nodes, BowVector doubles (bit for bit) and FeatureVector."""
import glob
import os
import tempfile

import numpy as np
import pytest

import bow_io
import oracle_lib as O
from ygz import Y

pytestmark = pytest.mark.gpu
GOLDEN = sorted(glob.glob(os.path.join(O.ROOT, "tests", "golden", "bow_*.npz")))
KEYS = ["word", "node", "weight", "bow_word", "bow_value", "fv_node", "fv_start", "fv_feat"]


def _same(a, b, ctx):
    for k in KEYS:
        x, y = np.asarray(a[k]), np.asarray(b[k])
        assert x.shape == y.shape and x.tobytes() == y.astype(x.dtype).tobytes(), (ctx, k)


@pytest.fixture(scope="module")
def matcher():
    m = Y.Matcher()
    yield m
    m.close()


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[4:-4] for p in GOLDEN])
def test_binary_loader_and_transform_vs_reference_golden(matcher, path):
    g = np.load(path)
    with tempfile.TemporaryDirectory() as td:
        vp = os.path.join(td, "voc.bin")
        open(vp, "wb").write(g["voc_bin"].tobytes())
        voc = Y.Vocabulary.load_binary(vp)
    a = voc.arrays()
    assert voc.n_nodes == len(g["tree_parent"]) and voc.k == int(g["tree_k"]) and voc.L == int(g["tree_L"])
    assert np.array_equal(a["child_start"], g["tree_child_start"]) and np.array_equal(a["child_list"], g["tree_child_list"])
    assert np.array_equal(a["word_id"], g["tree_word_id"]) and a["weight"].tobytes() == g["tree_weight"].tobytes()
    frames = [g["f%d_desc" % i] for i in range(int(g["n_frames"]))]          # ragged batch incl. 1 and 0 features
    got = Y.bow_transform(matcher, voc, frames, int(g["levelsup"]))
    for i in range(len(frames)):
        _same(got[i], {k: g["f%d_%s" % (i, k)] for k in KEYS}, (path, i))
    voc.close()


def test_text_loader_matches_binary(matcher):
    """ORBvoc.txt format (TemplatedVocabulary::loadFromTextFile): header 'k L scoring weighting', then one line per node
    'parent isLeaf d0 .. d31 weight' in NodeId order."""
    g = np.load(GOLDEN[0])
    n = len(g["tree_parent"]) - 1                       # without the duplicate node of the binary loader's read loop
    with tempfile.TemporaryDirectory() as td:
        tp = os.path.join(td, "voc.txt")
        with open(tp, "w") as f:
            f.write("%d %d 0 0\n" % (int(g["tree_k"]), int(g["tree_L"])))
            for i in range(1, n):
                leaf = g["tree_child_start"][i] == g["tree_child_start"][i + 1]
                f.write("%d %d %s %r\n" % (g["tree_parent"][i], int(leaf), " ".join(str(int(b)) for b in g["tree_desc"][i]),
                                           float(g["tree_weight"][i])))
            f.write("\n")
        voc = Y.Vocabulary.load_text(tp)
    a = voc.arrays()
    assert voc.n_nodes == n
    assert np.array_equal(a["desc"][1:], g["tree_desc"][1:n]) and a["weight"].tobytes() == g["tree_weight"][:n].tobytes()
    assert np.array_equal(a["word_id"], g["tree_word_id"][:n])
    got = Y.bow_transform(matcher, voc, [g["f0_desc"]], int(g["levelsup"]))[0]
    _same(got, {k: g["f0_%s" % k] for k in KEYS}, "text")
    voc.close()


@pytest.mark.parametrize("k,L,levelsup,p_stop,nfeat,seed", [(10, 5, 4, 0.0, (1000, 1000, 977, 3), 1), (10, 6, 4, 0.3, (2000, 1500), 2),
                                                             (3, 8, 2, 0.1, (4096,), 3), (17, 2, 5, 0.0, (600, 0, 1), 4)])
def test_transform_vs_oracle_synthetic_trees(matcher, k, L, levelsup, p_stop, nfeat, seed):
    t = bow_io.random_tree(k, L, seed, p_stop=p_stop)
    ovoc = O.Vocabulary.from_arrays(t["k"], t["L"], t["child_start"], t["child_list"], t["desc"], t["word_id"], t["weight"])
    voc = Y.Vocabulary.from_arrays(t["k"], t["L"], t["child_start"], t["child_list"], t["desc"], t["word_id"], t["weight"])
    frames = [bow_io.features_near_nodes(t, n, seed * 10 + i) if n else np.zeros((0, 32), np.uint8) for i, n in enumerate(nfeat)]
    got = Y.bow_transform(matcher, voc, frames, levelsup)
    for i, f in enumerate(frames):
        want = ovoc.transform(f, levelsup)
        _same(got[i], want, (k, L, i))
        if len(f) > 100:
            assert len(want["bow_word"]) > 10
    voc.close(); ovoc.close()


def test_feature_vector_feeds_triangulation_layout(matcher):
    """fv_node / fv_start / fv_feat are the CSR arrays ygzorb_keyframe_features takes: ascending ids, every live feature once."""
    t = bow_io.random_tree(10, 4, 9)
    voc = Y.Vocabulary.from_arrays(t["k"], t["L"], t["child_start"], t["child_list"], t["desc"], t["word_id"], t["weight"])
    f = bow_io.features_near_nodes(t, 1500, 10)
    r = Y.bow_transform(matcher, voc, [f], 2)[0]
    assert np.all(np.diff(r["fv_node"]) > 0) and r["fv_start"][0] == 0 and np.all(np.diff(r["fv_start"]) > 0)
    assert sorted(r["fv_feat"].tolist()) == np.nonzero(r["weight"] > 0)[0].tolist()
    for j in range(len(r["fv_node"])):
        seg = r["fv_feat"][r["fv_start"][j]:r["fv_start"][j + 1]]
        assert np.all(np.diff(seg) > 0) and np.all(r["node"][seg] == r["fv_node"][j])
    assert abs(r["bow_value"].sum() - 1.0) < 1e-12
    voc.close()


def test_vocabulary_rejects_malformed_trees():
    t = bow_io.random_tree(4, 2, 1)
    bad = t["child_list"].copy(); bad[0] = 0                               # a cycle through the root
    with pytest.raises(RuntimeError):
        Y.Vocabulary.from_arrays(4, 2, t["child_start"], bad, t["desc"], t["word_id"], t["weight"])
    bad = t["child_list"].copy(); bad[1] = len(t["word_id"])               # out of range
    with pytest.raises(RuntimeError):
        Y.Vocabulary.from_arrays(4, 2, t["child_start"], bad, t["desc"], t["word_id"], t["weight"])
