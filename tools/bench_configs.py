"""Secondary benchmarks for the other BASELINE.json configs (bench.py covers configs[1]):
  --config 0 : one 640x480 frame, 1000 features — single-frame latency through the C ABI (host in/out)
  --config 1 : 752x480 stream with the reference's own EuRoC.yaml extractor parameters (scale 2.0, 4 levels), device-resident batch
  --config 2 : KITTI-shape 1241x376 stereo pair, 2000 features/image + ComputeStereoMatches (host in/out)
  --config 3 : batched extraction of N_FRAMES 1280x720 frames, 2000 features, frame-sharded over the ranks
  --config 4 : brute-force kNN-2 + ratio, NQ x NT 256-bit rows, query-sharded over the ranks
Run under torchrun for > 1 GPU.  One JSON line per run on rank 0; times are CUDA-event device times, max over ranks
(configs 0/2: host wall clock around synchronous calls, since the host round trip is the point)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb-ygz-slam_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import sharding  # noqa: E402
import synth  # noqa: E402
import ygzorb_ctypes as Y  # noqa: E402


def setup():
    rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    return rank, local, world, dev


_KEEP = []   # handles and pinned buffers stay alive until the orderly shutdown at the end of main()


def timed_gather(fn, dev):
    """Device time of the result gather (the only communication of the sharded configs), max over ranks.  One small
    collective first: NCCL builds its communicator lazily, which is set-up, not gather time."""
    warm = torch.zeros((4, 8), dtype=torch.int32, device=dev)
    sharding.gather_rows(warm[: 4 // dist.get_world_size() or 1], (4 // dist.get_world_size() or 1) * dist.get_world_size())
    torch.cuda.synchronize()
    dist.barrier()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record()
    out = fn()
    g1.record()
    torch.cuda.synchronize()
    return out, sharding.max_over_ranks(g0.elapsed_time(g1), dev)


def single_frame(args, rank, local, world, dev):
    img = synth.textured_frame(640, 480, 7)
    ex = Y.Extractor(1000, device=local)
    pinned = torch.from_numpy(img).pin_memory()
    cap = ex.max_keypoints(640, 480)
    counts = torch.zeros(1, dtype=torch.int32).pin_memory(); kps = torch.zeros((cap, 7)).pin_memory()
    desc = torch.zeros((cap, 32), dtype=torch.uint8).pin_memory()
    def once():
        ex.extract_raw(pinned.data_ptr(), 1, 640, 480, 640, 640 * 480, 0, cap, counts.data_ptr(), kps.data_ptr(), desc.data_ptr())
    for _ in range(20):
        once()
    t0 = time.perf_counter()
    for _ in range(args.reps):
        once()
    dt = (time.perf_counter() - t0) / args.reps
    return {"config": "one 640x480 frame, 1000 features, host in/out, synchronous call", "latency_ms": dt * 1e3,
            "frames_per_s": 1 / dt, "keypoints": int(counts[0]), "launches_per_frame": 11}


def stereo(args, rank, local, world, dev):
    """Two legs: (a) the numpy binding, Left then Right then the match, three synchronous calls; (b) what Frame's constructor does
    in the reference — Left and Right extractors at the same time (src/Frame.cc:728-731: two threads there, two CUDA streams
    here), keypoints / descriptors kept on the device for ComputeStereoMatches and copied to pinned host memory beside it."""
    import ctypes as C
    w, h, nf = 1241, 376, 2000
    left, right = synth.stereo_pair(w, h, 6)
    exl, exr, mt = Y.Extractor(nf, device=local), Y.Extractor(nf, device=local), Y.Matcher(local)
    bf, b = 386.1448, 386.1448 / 718.856
    def once():
        kl, dl = exl.extract(left); kr, dr = exr.extract(right)
        return mt.stereo_match(exl, exr, 0, kl, dl, kr, dr, bf, b), len(kl), len(kr)
    for _ in range(5):
        once()
    t0 = time.perf_counter()
    for _ in range(args.reps):
        (ur, dp), nl, nr = once()
    dt = (time.perf_counter() - t0) / args.reps
    # ---- (b) concurrent, device-resident intermediate results
    cap = exl.max_keypoints(w, h)
    pin = [torch.from_numpy(left).pin_memory(), torch.from_numpy(right).pin_memory()]
    d_cnt = [torch.zeros(1, dtype=torch.int32, device=dev) for _ in range(2)]
    d_kps = [torch.zeros((cap, 7), device=dev) for _ in range(2)]
    d_desc = [torch.zeros((cap, 32), dtype=torch.uint8, device=dev) for _ in range(2)]
    h_cnt = [torch.zeros(1, dtype=torch.int32).pin_memory() for _ in range(2)]
    h_kps = [torch.zeros((cap, 7)).pin_memory() for _ in range(2)]
    h_desc = [torch.zeros((cap, 32), dtype=torch.uint8).pin_memory() for _ in range(2)]
    h_ur, h_dp = torch.zeros(cap).pin_memory(), torch.zeros(cap).pin_memory()
    streams = [torch.cuda.ExternalStream(e.stream(), device=dev) for e in (exl, exr)]
    _KEEP.extend([exl, exr, mt, pin, h_cnt, h_kps, h_desc, h_ur, h_dp, streams])
    L = Y.lib()
    def once2():
        for i, e in enumerate((exl, exr)):
            e.extract_raw(pin[i].data_ptr(), 1, w, h, w, w * h, Y.OUT_DEVICE | Y.NO_SYNC, cap, d_cnt[i].data_ptr(), d_kps[i].data_ptr(),
                          d_desc[i].data_ptr())
            with torch.cuda.stream(streams[i]):
                h_cnt[i].copy_(d_cnt[i], non_blocking=True)
                h_kps[i].copy_(d_kps[i], non_blocking=True); h_desc[i].copy_(d_desc[i], non_blocking=True)
        exl.sync(); exr.sync()
        n_l, n_r = int(h_cnt[0][0]), int(h_cnt[1][0])
        Y.check(L.ygzorb_stereo_match(mt.h, exl.h, exr.h, 0, C.c_void_p(d_kps[0].data_ptr()), C.c_void_p(d_desc[0].data_ptr()), n_l,
                                      C.c_void_p(d_kps[1].data_ptr()), C.c_void_p(d_desc[1].data_ptr()), n_r, C.c_float(bf), C.c_float(b),
                                      Y.IN_DEVICE, C.c_void_p(h_ur.data_ptr()), C.c_void_p(h_dp.data_ptr())))
        return n_l
    for _ in range(5):
        n_l = once2()
    assert n_l == nl and np.array_equal(h_ur[:n_l].numpy(), ur) and np.array_equal(h_dp[:n_l].numpy(), dp), "the two legs disagree"
    t0 = time.perf_counter()
    for _ in range(args.reps):
        once2()
    dt2 = (time.perf_counter() - t0) / args.reps
    return {"config": "1241x376 stereo pair, 2000 features/image + ComputeStereoMatches, host in/out",
            "latency_ms_sequential_numpy_binding": dt * 1e3, "latency_ms": dt2 * 1e3, "pairs_per_s": 1 / dt2,
            "note": "latency_ms: Left/Right extractors concurrently on their own streams, keypoints/descriptors device-resident for the match, pinned host buffers",
            "left_kp": nl, "right_kp": nr, "stereo_matches": int((ur >= 0).sum())}


def batched(args, rank, local, world, dev):
    w, h, nf, n_total, chunk = 1280, 720, 2000, args.frames, args.chunk
    b, e = sharding.shard_range(n_total, rank, world)
    n = e - b
    base = synth.frame_batch(w, h, min(chunk, max(n, 1)), seed0=200 + rank, distinct=8)
    d_in = torch.from_numpy(base).to(dev)                     # the shard is streamed chunk by chunk from HBM
    ex = Y.Extractor(nf, max_batch=chunk, device=local)
    cap = ex.max_keypoints(w, h)
    counts = torch.zeros(max(n, 1), dtype=torch.int32, device=dev)
    kps = torch.zeros((max(n, 1), cap, 7), device=dev); desc = torch.zeros((max(n, 1), cap, 32), dtype=torch.uint8, device=dev)
    stream = torch.cuda.ExternalStream(ex.stream(), device=dev)
    flags = Y.IN_DEVICE | Y.OUT_DEVICE | Y.NO_SYNC
    def run():
        for f0 in range(0, n, chunk):
            m = min(chunk, n - f0)
            ex.extract_raw(d_in.data_ptr(), m, w, h, w, w * h, flags, cap, counts[f0:].data_ptr(), kps[f0:].data_ptr(), desc[f0:].data_ptr())
    run(); ex.sync()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream); run(); e1.record(stream); ex.sync()
    ms = e0.elapsed_time(e1)
    gather_ms = 0.0
    if world > 1:   # the only communication: gather of (count, keypoints, descriptors) per frame to rank 0
        _, gather_ms = timed_gather(lambda: (sharding.gather_rows(counts[:n].unsqueeze(1), n_total), sharding.gather_rows(kps[:n], n_total),
                                             sharding.gather_rows(desc[:n], n_total)), dev)
        ms = sharding.max_over_ranks(ms, dev)
    staged = 13392083
    return {"config": "%d frames 1280x720, 2000 features, frame-sharded over %d GPU(s), chunk %d (the chunk's frames are reused from HBM)"
                      % (n_total, world, chunk), "ms": ms, "frames_per_s": n_total / (ms * 1e-3), "gather_ms": gather_ms,
            "frames_per_s_incl_gather": n_total / ((ms + gather_ms) * 1e-3),
            "keypoints_per_frame": float(counts[:n].float().mean().item()) if n else 0,
            "staged_GBps": staged * n_total / (ms * 1e-3) / 1e9, "scaling": "strong"}


def euroc_yaml(args, rank, local, world, dev):
    """Secondary line of configs[1]: the extractor parameters of the reference's own
    Examples/Monocular/EuRoC.yaml:35-38 (scaleFactor 2.0, 4 levels; level 3 has no FAST cell row and yields nothing), 752x480,
    1000 features, a device-resident batch of 512 frames."""
    w, h, nf, n = 752, 480, 1000, 512
    base = synth.frame_batch(w, h, n, seed0=300 + rank, distinct=8)
    d_in = torch.from_numpy(base).to(dev)
    ex = Y.Extractor(nf, 2.0, 4, 20, 7, max_batch=n, device=local)
    cap = ex.max_keypoints(w, h)
    counts = torch.zeros(n, dtype=torch.int32, device=dev)
    kps = torch.zeros((n, cap, 7), device=dev); desc = torch.zeros((n, cap, 32), dtype=torch.uint8, device=dev)
    stream = torch.cuda.ExternalStream(ex.stream(), device=dev)
    flags = Y.IN_DEVICE | Y.OUT_DEVICE | Y.NO_SYNC
    _KEEP.extend([ex, stream])
    def run():
        ex.extract_raw(d_in.data_ptr(), n, w, h, w, w * h, flags, cap, counts.data_ptr(), kps.data_ptr(), desc.data_ptr())
    for _ in range(3):
        run()
    ex.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(10):
        run()
    e1.record(stream); ex.sync()
    ms = e0.elapsed_time(e1) / 10
    return {"config": "752x480, 1000 features, scaleFactor 2.0 / 4 levels (Examples/Monocular/EuRoC.yaml), 512 device-resident frames per step",
            "ms_per_step": ms, "frames_per_s": n / (ms * 1e-3), "keypoints_per_frame": float(counts.float().mean().item())}


def knn(args, rank, local, world, dev):
    nq, nt = args.nq, args.nt
    b, e = sharding.shard_range(nq, rank, world)
    rng = np.random.default_rng(5)
    train = torch.from_numpy(rng.integers(0, 256, size=(nt, 32), dtype=np.uint8)).to(dev) if rank == 0 else torch.zeros((nt, 32), dtype=torch.uint8, device=dev)
    if world > 1:
        dist.broadcast(train, src=0)                          # train rows replicated once (32 MB for 1M)
    g = torch.Generator(device=dev); g.manual_seed(100 + rank)
    n = e - b
    query = torch.randint(0, 256, (max(n, 1), 32), dtype=torch.uint8, device=dev, generator=g)
    half = n // 2
    if half:
        src = torch.randint(0, nt, (half,), device=dev, generator=g)
        noise = (torch.rand((half, 32, 8), device=dev, generator=g) < 0.1).to(torch.uint8)
        bits = (noise << torch.arange(8, device=dev, dtype=torch.uint8)).sum(dim=2).to(torch.uint8)
        query[:half] = train[src] ^ bits
    mt = Y.Matcher(local)
    out = torch.zeros((max(n, 1), 4), dtype=torch.int32, device=dev)
    cols = [torch.zeros(max(n, 1), dtype=torch.int32, device=dev) for _ in range(4)]
    stream = torch.cuda.ExternalStream(mt.stream(), device=dev)
    flags = Y.IN_DEVICE | Y.OUT_DEVICE | Y.NO_SYNC
    def run():
        mt.hamming_knn2_raw(query.data_ptr(), n, train.data_ptr(), nt, args.ratio, 50, flags, *[c.data_ptr() for c in cols])
    mt.hamming_knn2_raw(query.data_ptr(), min(n, 4096), train.data_ptr(), nt, args.ratio, 50, flags, *[c.data_ptr() for c in cols]); mt.sync()
    peak = mt.popc_peak()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream); run(); e1.record(stream); mt.sync()
    ms = e0.elapsed_time(e1)
    out = torch.stack(cols, dim=1)
    gather_ms = 0.0
    if world > 1:
        full, gather_ms = timed_gather(lambda: sharding.gather_rows(out[:n], nq), dev)
        ms = sharding.max_over_ranks(ms, dev)
        matches = int((full[:, 3] >= 0).sum().item()) if rank == 0 else 0
    else:
        matches = int((out[:n, 3] >= 0).sum().item())
    pairs = nq * nt / (ms * 1e-3)
    return {"config": "brute-force Hamming kNN k=2 + ratio %.2f, %d x %d rows, query-sharded over %d GPU(s)" % (args.ratio, nq, nt, world),
            "ms": ms, "pairs_per_s": pairs, "queries_per_s": nq / (ms * 1e-3), "gather_ms": gather_ms, "matches": matches,
            "popc_peak_per_s_per_gpu": peak, "frac_of_popc_peak": pairs * 8 / (peak * world), "scaling": "strong"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, required=True)
    ap.add_argument("--reps", type=int, default=200)
    ap.add_argument("--frames", type=int, default=4096)
    ap.add_argument("--chunk", type=int, default=256)
    ap.add_argument("--nq", type=int, default=1 << 20)
    ap.add_argument("--nt", type=int, default=1 << 20)
    ap.add_argument("--ratio", type=float, default=0.6)
    args = ap.parse_args()
    rank, local, world, dev = setup()
    fn = {0: single_frame, 1: euroc_yaml, 2: stereo, 3: batched, 4: knn}[args.config]
    res = fn(args, rank, local, world, dev)
    res["n_gpus"] = world
    if rank == 0:
        print(json.dumps(res), flush=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.destroy_process_group()
    sys.stdout.flush(); sys.stderr.flush()
    torch.cuda.synchronize()
    Y.close_all()   # orderly shutdown: handles first, then a normal interpreter exit (exit hooks run)


if __name__ == "__main__":
    main()
