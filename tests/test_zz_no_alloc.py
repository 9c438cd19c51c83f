"""ygzorb_hamming_knn2_noalloc (include/ygzorb.h): the kNN-2 + ratio kernels with their scratch in a static __device__
array instead of the matcher handle's cudaMalloc'd pool.  It has no counterpart in the reference;
its results must equal those of the existing entry point ygzorb_hamming_knn2 — integer outputs,
compared for equality — on two calls in a row (the array keeps the previous call's contents).  Collected last on
purpose."""
import numpy as np
import pytest

import synth
from ygz import Y

pytestmark = pytest.mark.gpu


def _both(torch, q, t, ratio, th, stream=None):
    nq, nt = len(q), len(t)
    want = Y.hamming_knn2(q, t, ratio, th)                        # existing entry point (host buffers, handle-owned scratch)
    dq = torch.from_numpy(q).cuda()
    dt = torch.from_numpy(t).cuda() if nt else torch.zeros(32, dtype=torch.uint8, device="cuda")
    out = [torch.full((max(nq, 1),), -7, dtype=torch.int32, device="cuda") for _ in range(4)]
    torch.cuda.synchronize()
    Y.hamming_knn2_noalloc(0, stream, dq.data_ptr(), nq, dt.data_ptr() if nt else None, nt, ratio, th,
                           *[o.data_ptr() for o in out])
    torch.cuda.synchronize()
    got = [o.cpu().numpy()[:nq] for o in out]
    return want, got


def test_noalloc_matches_existing_entry_point_twice():
    import torch
    # first call: many queries, one train split; second call: few queries -> many train splits, i.e. the partial-result
    # layout in the shared scratch array differs from what the first call left behind
    cases = [(5000, 20000, 0.6, 50, 1), (300, 70001, 0.9, 100, 2), (5000, 20000, 0.75, 64, 3)]
    for nq, nt, ratio, th, seed in cases:
        q, t = synth.knn_case(nq, nt, seed)
        want, got = _both(torch, q, t, ratio, th)
        for w, g in zip(want, got):
            assert np.array_equal(w, g)
        assert (got[3] >= 0).sum() > 0


def test_noalloc_on_a_side_stream_and_empty_train():
    import torch
    s = torch.cuda.Stream()
    q, t = synth.knn_case(1000, 4096, 5)
    want, got = _both(torch, q, t, 0.8, 100, stream=s.cuda_stream)
    for w, g in zip(want, got):
        assert np.array_equal(w, g)
    want, got = _both(torch, q, t[:0], 0.8, 100)
    for w, g in zip(want, got):
        assert np.array_equal(w, g)
    assert (got[0] == -1).all() and (got[1] == 256).all()


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    import torch
    dq = torch.zeros(64, dtype=torch.uint8, device="cuda")        # never read: the size check comes first
    nq = Y.NOALLOC_SCRATCH_BYTES // 12 + 1
    with pytest.raises(Y.YgzError, match="static array"):
        Y.hamming_knn2_noalloc(0, None, dq.data_ptr(), nq, dq.data_ptr(), 2, 0.6, 50, None, None, None, None)
    assert Y.NOALLOC_MAX_QUERIES * 12 <= Y.NOALLOC_SCRATCH_BYTES
