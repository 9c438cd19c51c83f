"""Brute-force kNN-2 throughput (BASELINE.json config 5 shape, bounded) + measured POPC peak."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import synth  # noqa: E402
from ygz import Y  # noqa: E402


def main():
    nq = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
    nt = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 20
    m = Y.Matcher()
    peak = m.popc_peak()
    print("popc peak %.3f Tpopc/s" % (peak / 1e12))
    q, t = synth.knn_case(nq, nt, 1)
    dq, dt = torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda()
    out = [torch.zeros(nq, dtype=torch.int32, device="cuda") for _ in range(4)]
    stream = torch.cuda.ExternalStream(m.stream())
    for variant in (0, 1, 2, 3):
        m.set_knn_variant(variant)
        for _ in range(2):
            m.hamming_knn2_raw(dq.data_ptr(), nq, dt.data_ptr(), nt, 0.6, 50, Y.IN_DEVICE | Y.OUT_DEVICE, *[o.data_ptr() for o in out])
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        reps = 3
        for _ in range(reps):
            m.hamming_knn2_raw(dq.data_ptr(), nq, dt.data_ptr(), nt, 0.6, 50, Y.IN_DEVICE | Y.OUT_DEVICE | Y.NO_SYNC,
                               *[o.data_ptr() for o in out])
        e1.record(stream)
        m.sync()
        ms = e0.elapsed_time(e1) / reps
        pairs = nq * nt / (ms * 1e-3)
        print("variant %d: %.3f ms, %.1f Gpairs/s, %.3f of popc peak (8 popc/pair)" % (variant, ms, pairs / 1e9, pairs * 8 / peak),
              "matches", int((out[3] >= 0).sum().item()))


if __name__ == "__main__":
    main()
