"""The multi-GPU layer (orb-ygz-slam_b200/csrc/comm.cu) on ONE GPU: with world == 1 the sharded calls must return exactly what the
single-GPU entry points return — the frame-sharded extraction packs the records (counts first, then the ragged payload, chunk by
chunk) and the query-sharded kNN writes its slice in place.  The N > 1 path is exercised by bench.py under torchrun (its `sharded`
block) and by tests/test_sharding_gloo.py for the layout."""
import numpy as np
import pytest

import synth
from ygz import Y

pytestmark = pytest.mark.gpu


def test_extract_batch_sharded_world1_equals_plain():
    import torch
    w, h, nf, n = 640, 480, 1000, 37
    frames = synth.frame_batch(w, h, n, seed0=50, distinct=5)
    frames[3] = 128                                           # a frame without a single keypoint: an empty record range
    ex = Y.Extractor(nf, max_batch=16)
    want = ex.extract_batch(frames)
    cap = ex.max_keypoints(w, h)
    comm = Y.Comm()
    d_in = torch.from_numpy(frames).cuda()
    counts = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    off = torch.full((n,), -1, dtype=torch.int64, device="cuda")
    kps = torch.zeros((n * cap, 7), dtype=torch.float32, device="cuda")
    desc = torch.zeros((n * cap, 32), dtype=torch.uint8, device="cuda")
    for chunk, in_dev in [(16, True), (5, True), (16, False)]:
        counts.fill_(-1); off.fill_(-1)
        src = d_in.data_ptr() if in_dev else frames.ctypes.data
        comm.extract_batch_sharded(ex, src, n, w, h, w, w * h, Y.IN_DEVICE if in_dev else 0, cap, chunk, counts.data_ptr(), off.data_ptr(),
                                   kps.data_ptr(), desc.data_ptr())
        c, o = counts.cpu().numpy(), off.cpu().numpy()
        k = kps.cpu().numpy().view(Y.KP_DTYPE).reshape(-1); d = desc.cpu().numpy()
        assert np.array_equal(c, [len(x[0]) for x in want]) and c[3] == 0
        assert np.array_equal(o, np.concatenate([[0], np.cumsum(c)[:-1]]))          # one rank: one region, fully packed
        for f in range(n):
            assert k[o[f]:o[f] + c[f]].tobytes() == want[f][0].tobytes() and np.array_equal(d[o[f]:o[f] + c[f]], want[f][1])
    assert comm.launch_count() > 0
    comm.close(); ex.close()


def test_knn2_sharded_world1_equals_plain():
    import torch
    q, t = synth.knn_case(3001, 5000, 9)
    want = Y.hamming_knn2(q, t, 0.7, 60)
    m, comm = Y.Matcher(), Y.Comm()
    dq, dt = torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda()
    out = [torch.full((len(q),), -9, dtype=torch.int32, device="cuda") for _ in range(4)]
    comm.hamming_knn2_sharded(m, dq.data_ptr(), len(q), dt.data_ptr(), len(t), 0.7, 60, Y.BROADCAST_TRAIN, *[o.data_ptr() for o in out])
    for a, b in zip(want, out):
        assert np.array_equal(a, b.cpu().numpy())
    comm.close(); m.close()


def test_shard_range_abi():
    for n in (0, 1, 7, 4096, 1000003):
        for world in (1, 2, 3, 8):
            r = [Y.shard_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n and all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            sizes = [e - b for b, e in r]
            assert max(sizes) - min(sizes) <= 1
