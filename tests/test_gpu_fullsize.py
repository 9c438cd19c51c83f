"""Full-size parity checks through size-independent properties (BASELINE.json's configurations are too large for the CPU oracle
to finish in seconds, so the oracle anchors a few frames / rows and properties carry the result to the whole batch):

  * configs[1] (512 frames 752x480 per step, the bench's batch): the batch is eight distinct frames repeated in a shuffled order —
    every copy must give byte-identical keypoints and descriptors whatever its position in the batch (independence of batch
    position / chunk), the eight originals equal the oracle, and a checksum over the whole result does not depend on how the
    library chunks the batch;
  * SearchByProjection on the same batch: a frame's assignments do not depend on its position in the batch, and equal the oracle;
  * configs[3] shape (1280x720, 2000 features): the same on 64 frames;
  * configs[4] shape (brute-force kNN-2): 65,536 queries x 1,048,576 train rows — queries that are copies of train rows find their row
    at distance 0; d1 <= d2; a sample of queries equals numpy's exhaustive answer; the result does not depend on the split count
    (ygzorb_hamming_knn2 vs ygzorb_hamming_knn2_noalloc, which plans its splits for an 8 MiB scratch)."""
import hashlib

import numpy as np
import pytest
import torch

import oracle_lib as O
import synth
from ygz import Y

pytestmark = pytest.mark.gpu


def _digest(outs):
    h = hashlib.sha1()
    for kps, desc in outs:
        h.update(np.int32(len(kps)).tobytes()); h.update(kps.tobytes()); h.update(desc.tobytes())
    return h.hexdigest()


def _batch(w, h, n, distinct, seed):
    base = [synth.textured_frame(w, h, seed + i) if i % 3 else synth.mixed_frame(w, h, seed + i) for i in range(distinct)]
    order = np.random.default_rng(seed).integers(0, distinct, n)
    order[:distinct] = np.arange(distinct)          # every original appears at least once, up front
    return np.stack([base[k] for k in order]), order, base


@pytest.mark.parametrize("w,h,nf,n,distinct", [(752, 480, 1000, 512, 8), (1280, 720, 2000, 64, 4)])
def test_extraction_is_independent_of_batch_position(w, h, nf, n, distinct):
    frames, order, base = _batch(w, h, n, distinct, 700)
    p = O.params(nf, 1.2, 8, 20, 7)
    want = [O.extract(p, b) for b in base]
    ex = Y.Extractor(nf, 1.2, 8, 20, 7, max_batch=n)
    outs = ex.extract_batch(frames)
    for f in range(n):
        k = order[f]
        assert outs[f][0].tobytes() == want[k][0].tobytes(), "frame %d (copy of %d)" % (f, k)
        assert np.array_equal(outs[f][1], want[k][1])
    whole = _digest(outs)
    for chunk in (96, 37):                          # a different chunking of the same host batch: same bytes out
        ex.set_pipeline(chunk)
        assert _digest(ex.extract_batch(frames)) == whole, "chunk %d" % chunk
    ex.close()


def test_search_by_projection_is_independent_of_batch_position():
    """512 frames x 3000 MapPoints (the bench's step): frames are four distinct (keypoints, MapPoints, claimed) sets repeated in a
    shuffled order; each copy's assignment equals the oracle's for its original."""
    w, h, n, distinct = 752, 480, 512, 4
    ex = Y.Extractor(1000)
    sf, _, _ = synth.scale_tables()
    sets = []
    for k in range(distinct):
        kps, desc = ex.extract(synth.textured_frame(w, h, 800 + k))
        mp, u_right, claimed = synth.projection_case(kps, desc, 3000, seed=900 + k)
        sets.append((kps, desc, None, claimed, mp))
    ex.close()
    bounds = (0.0, 0.0, float(w), float(h))
    want = [O.search_by_projection(kps, desc, None, claimed, bounds, sf, mp, 1.0, False, 0.8) for kps, desc, _, claimed, mp in sets]
    order = np.random.default_rng(5).integers(0, distinct, n)
    m = Y.Matcher()
    got = m.search_by_projection([sets[k][:4] for k in order], [sets[k][4] for k in order], bounds, sf, 1.0, False, 0.8)
    m.close()
    for f in range(n):
        k = order[f]
        assert got[f][0] == want[k][0] and want[k][0] > 500, f
        assert np.array_equal(got[f][1], want[k][1]) and np.array_equal(got[f][2], want[k][2]), f


def test_knn2_at_a_million_train_rows():
    nq, nt = 65536, 1 << 20
    rng = np.random.default_rng(11)
    train = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
    query = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    src = rng.integers(0, nt, nq // 2)
    query[: nq // 2] = train[src]                                   # exact copies: nearest neighbour at distance 0
    flip = rng.integers(0, 256, (nq // 4, 2))
    for i in range(nq // 4):                                        # a quarter of them with two flipped bits: distance 2
        query[i, flip[i, 0] % 32] ^= 1 << (flip[i, 1] % 8)
        query[i, (flip[i, 0] + 7) % 32] ^= 1 << ((flip[i, 1] + 3) % 8)
    dev = torch.device("cuda", 0)
    dq, dt = torch.from_numpy(query).to(dev), torch.from_numpy(train).to(dev)
    res = [torch.full((nq,), -7, dtype=torch.int32, device=dev) for _ in range(4)]
    m = Y.Matcher()
    m.hamming_knn2_raw(dq.data_ptr(), nq, dt.data_ptr(), nt, 0.6, 50, Y.IN_DEVICE | Y.OUT_DEVICE, *[r.data_ptr() for r in res])
    m.sync()
    idx, d1, d2, match = (r.cpu().numpy() for r in res)
    # properties over ALL queries
    assert (d1 <= d2).all() and (idx >= 0).all() and (idx < nt).all()
    assert (d1[nq // 4: nq // 2] == 0).all() and (d1[: nq // 4] == 2).all()
    same = train[idx[: nq // 2]]
    assert (np.unpackbits(same ^ query[: nq // 2], axis=1).sum(axis=1) == d1[: nq // 2]).all()   # the row it names is at d1
    ok = (d1 <= 50) & (d1.astype(np.float64) < 0.6 * d2.astype(np.float64).astype(np.float32))
    assert np.array_equal(match >= 0, (d1 <= 50) & (d1.astype(np.float32) < np.float32(0.6) * d2.astype(np.float32)))
    assert ok.sum() >= nq // 2 - 16
    # a sample against numpy's exhaustive search (first index wins on ties, as the serial loop would)
    for q in rng.integers(0, nq, 24):
        d = np.unpackbits(train ^ query[q], axis=1).sum(axis=1, dtype=np.int32)
        o = np.lexsort((np.arange(nt), d))[:2]
        assert idx[q] == o[0] and d1[q] == d[o[0]] and d2[q] == d[o[1]], q
    # the split count does not matter: the no-alloc entry point plans its splits for an 8 MiB scratch array
    res2 = [torch.full((nq,), -7, dtype=torch.int32, device=dev) for _ in range(4)]
    Y.hamming_knn2_noalloc(0, None, dq.data_ptr(), nq, dt.data_ptr(), nt, 0.6, 50, *[r.data_ptr() for r in res2])
    torch.cuda.synchronize()
    for a, b in zip(res, res2):
        assert torch.equal(a, b)
    m.close()
