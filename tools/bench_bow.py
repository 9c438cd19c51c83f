"""Bag-of-words assignment on the shape the system runs it: FRAMES frames x 1000 ORB descriptors
against an ORBvoc-shaped vocabulary (k=10, L=6, 1,111,111 nodes), levelsup=4 (src/Frame.cc:498).
GPU: ygzorb_bow_transform, device-resident (CUDA events) and host in/out; CPU: the reference's own DBoW2
(oracle/_ref/ygz_ref_bow, all host threads) on a bounded sample, whose outputs also check the GPU results at full tree
size.  One JSON line."""
import argparse
import json
import os
import re
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb-ygz-slam_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import bow_io  # noqa: E402
import synth  # noqa: E402
import ygzorb_ctypes as Y  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=512)
    ap.add_argument("--feats", type=int, default=1000)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--cpu-frames", type=int, default=64)
    args = ap.parse_args()
    F, N = args.frames, args.feats
    tree = bow_io.full_tree(10, 6, 1)
    rng = np.random.default_rng(2)
    leaves = np.nonzero(tree["word_id"] >= 0)[0]
    src = rng.choice(leaves, size=F * N)
    desc = synth.flip_bits(tree["desc"][src], rng.integers(0, 60, size=F * N), rng).reshape(F, N, 32)
    dev = torch.device("cuda", 0)
    with tempfile.TemporaryDirectory() as td:
        vp = os.path.join(td, "voc.bin")
        bow_io.write_binary_voc(vp, tree)
        t0 = time.time()
        voc = Y.Vocabulary.load_binary(vp)
        load_s = time.time() - t0
        mt = Y.Matcher(0)
        stream = torch.cuda.ExternalStream(mt.stream(), device=dev)
        d_desc = torch.from_numpy(desc).to(dev)
        d_n = torch.full((F,), N, dtype=torch.int32, device=dev)
        outs = {"word": torch.zeros((F, N), dtype=torch.int32, device=dev), "node": torch.zeros((F, N), dtype=torch.int32, device=dev),
                "weight": torch.zeros((F, N), dtype=torch.float64, device=dev), "bc": torch.zeros(F, dtype=torch.int32, device=dev),
                "bw": torch.zeros((F, N), dtype=torch.int32, device=dev), "bv": torch.zeros((F, N), dtype=torch.float64, device=dev),
                "fc": torch.zeros(F, dtype=torch.int32, device=dev), "fn": torch.zeros((F, N), dtype=torch.int32, device=dev),
                "fs": torch.zeros((F, N + 1), dtype=torch.int32, device=dev), "ff": torch.zeros((F, N), dtype=torch.int32, device=dev)}
        order = ["word", "node", "weight", "bc", "bw", "bv", "fc", "fn", "fs", "ff"]
        import ctypes as C

        def run(flags, dptr, nptr, optrs):
            Y.check(Y.lib().ygzorb_bow_transform(mt.h, voc.h, C.c_void_p(dptr), C.c_void_p(nptr), F, N, 4, flags,
                                                 *[C.c_void_p(p) for p in optrs]))
        dflags = Y.IN_DEVICE | Y.OUT_DEVICE | Y.NO_SYNC
        optrs = [outs[k].data_ptr() for k in order]
        for _ in range(3):
            run(dflags, d_desc.data_ptr(), d_n.data_ptr(), optrs)
        mt.sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(args.steps):
            run(dflags, d_desc.data_ptr(), d_n.data_ptr(), optrs)
        e1.record(stream)
        mt.sync()
        ms = e0.elapsed_time(e1) / args.steps
        # host in/out (numpy arrays, synchronous call): what ORBVocabulary::transform pays per batch
        h = {k: v.cpu().numpy().copy() for k, v in outs.items()}
        n_np = np.full(F, N, np.int32)
        hptrs = [h[k].ctypes.data for k in order]
        run(0, desc.ctypes.data, n_np.ctypes.data, hptrs)
        t0 = time.time()
        for _ in range(3):
            run(0, desc.ctypes.data, n_np.ctypes.data, hptrs)
        ms_host = (time.time() - t0) / 3 * 1e3
        # the reference's DBoW2 on a bounded sample, all host threads; its outputs check the GPU's at full tree size
        cpu = None
        if bow_io.have_ref():
            S = min(args.cpu_frames, F)
            threads = os.cpu_count() or 1
            res, out = bow_io.ref_transform(vp, [desc[i] for i in range(S)], 4, threads=threads, repeat=4)
            m = re.search(r"([0-9.]+) features/s", out)
            cpu = {"value": float(m.group(1)), "unit": "features/s", "cores": threads, "kind": "reference",
                   "sample": "%d frames x %d features x 4 repeats, Thirdparty/DBoW2 transform" % (S, N)}
            for i in range(S):
                nb, nn = int(h["bc"][i]), int(h["fc"][i])
                assert np.array_equal(h["word"][i], res[i]["word"]) and np.array_equal(h["node"][i], res[i]["node"]), i
                assert np.array_equal(h["bw"][i, :nb], res[i]["bow_word"]) and h["bv"][i, :nb].tobytes() == res[i]["bow_value"].tobytes(), i
                assert np.array_equal(h["fn"][i, :nn], res[i]["fv_node"]) and np.array_equal(h["ff"][i, :h["fs"][i, nn]], res[i]["fv_feat"]), i
    line = {"metric": "bag-of-words assignment, features/s", "value": F * N / (ms * 1e-3), "unit": "features/s", "ms_per_step": ms,
            "frames_per_s": F / (ms * 1e-3), "e2e": {"value": F * N / (ms_host * 1e-3), "unit": "features/s", "ms_per_step": ms_host},
            "config": {"workload": "%d frames x %d descriptors, vocabulary k=10 L=6 (%d nodes, %.0f MB of node descriptors), levelsup 4"
                                   % (F, N, voc.n_nodes, voc.n_nodes * 32 / 1e6), "vocabulary_load_s": load_s},
            "cpu_baseline": cpu, "parity_vs_reference_sample": "bit-exact" if cpu else "not checked (no oracle/_ref)",
            "hamming_pairs_per_s": F * N * 60 / (ms * 1e-3)}
    print(json.dumps(line))
    torch.cuda.synchronize()
    Y.close_all()   # orderly shutdown: handles first, then a normal interpreter exit (exit hooks run)


if __name__ == "__main__":
    main()
