"""bench.py — headline benchmark of the ORB front-end hot path on B200 (contract: see DESIGN.md §Measurement).

A "step" is one pass of the hot path over one batch of synthetic EuRoC-shape frames
(BASELINE.json configs[1]: 752x480, 1000 features, scale 1.2, 8 levels, FAST 20/7): ORB extraction
of every frame, plus — once the matcher kernels are built — SearchByProjection of 3000 projected
MapPoints per frame.  `value` = frames/s with inputs resident in HBM; `e2e` = the same through the
C ABI with host (pinned) buffers, H2D and D2H inside the timed region.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--batch B]
Under torchrun (N > 1) every rank runs its own stream of frames (weak scaling, no data-path
collective: frames are independent); times are device times, max over ranks.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "orb-ygz-slam_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "ORB frames/sec (752x480, 1000 kp, 8 lvl) and Hamming matches/sec vs CPU"
W, H, NFEAT, SCALE, LEVELS, INI_TH, MIN_TH = 752, 480, 1000, 1.2, 8, 20, 7
N_MAPPOINTS = 3000
N_SLOTS = 2   # device buffer sets of the e2e leg (MapPoints in, results out); 3 measured 2 % slower
# algorithmic bytes per 752x480 frame, stage by stage (inputs read once + outputs written once)
STAGE_BYTES = {"pyramid": 1845634, "fast": 1117367, "blur": 2234734, "quadtree_orient_desc": 60000}
STAGED_TOTAL = 5257735
STAGE_NAMES = ["upload", "pyramid", "fast", "quadtree", "blur", "orient_desc", "download"]


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def bind_to_gpu_cpus(index):
    """Pins this rank to the CPUs NVML names as local to its GPU, BEFORE the pinned staging buffers are allocated: on a
    multi-socket box first-touch then places them on the GPU's own NUMA node, so N ranks do not all stream their frames out
    of one socket's memory.  Best effort (a container may hide the topology); returns what was done for the JSON line."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = [64 * i + b for i, wd in enumerate(words) for b in range(64) if (wd >> b) & 1]
        allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
        if allowed:
            os.sched_setaffinity(0, allowed)
            return {"cpus": len(allowed), "first": allowed[0]}
    except Exception as e:   # noqa: BLE001
        return {"error": str(e)[:80]}
    return None


class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.dev = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.dev, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                 "hw_power_brake_slowdown": 0x80}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.dev, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.dev)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.02)

    def start(self):
        if self.nv:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()

    def stop(self):
        self._stop.set()
        if self._thread:
            self._thread.join()
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def make_frames(n, seed0):
    import synth
    return synth.frame_batch(W, H, n, seed0=seed0, distinct=8)


def make_mixed_frames(n, seed0):
    import synth
    def one(seed):   # left half full contrast, right half at 0.18 of it (iniThFAST 20 finds nothing there, minThFAST 7 does), a flat band
        a = synth.textured_frame(W, H, seed).astype(np.float64)
        a[:, W // 2:] = 128 + (a[:, W // 2:] - 128) * 0.18
        a[H // 3: H // 3 + 45, :] = 97
        return np.clip(np.rint(a), 0, 255).astype(np.uint8)
    base = [one(seed0 + i) for i in range(min(8, n))]
    out = np.empty((n, H, W), np.uint8)
    for i in range(n):
        out[i] = np.roll(base[i % len(base)], shift=3 * (i // len(base)), axis=0)     # vertical roll: the halves stay halves
    return out


def retry_cell_fraction(img):
    """Share of the FAST cells of one frame whose iniThFAST pass comes back empty, i.e. which the reference detects a second
    time at minThFAST (src/ORBextractor.cc:764-770); counted with the oracle's pyramid and OpenCV's own detector on the
    reference's cell grid (:728-762).  Set-up only."""
    import cv2
    import oracle_lib as O
    import math
    p = O.params(NFEAT, SCALE, LEVELS, INI_TH, MIN_TH)
    ini, mn = cv2.FastFeatureDetector_create(INI_TH, True), cv2.FastFeatureDetector_create(MIN_TH, True)
    cells = retry = empty = 0
    for lv in O.pyramid(p, img):
        h, w = lv.shape
        max_bx, max_by = w - 16, h - 16
        width, height = float(max_bx - 16), float(max_by - 16)
        ncols, nrows = int(width / 30), int(height / 30)
        if ncols <= 0 or nrows <= 0:
            continue
        wc, hc = math.ceil(width / ncols), math.ceil(height / nrows)
        for i in range(nrows):
            iy = 16 + i * hc
            if iy >= max_by - 3:
                continue
            my = min(iy + hc + 6, max_by)
            for j in range(ncols):
                ix = 16 + j * wc
                if ix >= max_bx - 6:
                    continue
                mx = min(ix + wc + 6, max_bx)
                sub = np.ascontiguousarray(lv[iy:my, ix:mx])
                cells += 1
                if not ini.detect(sub):
                    retry += 1
                    if not mn.detect(sub):
                        empty += 1
    return {"fast_cells_per_frame": cells, "retry_cell_fraction": retry / max(cells, 1), "empty_cell_fraction": empty / max(cells, 1)}


def ref_binary():
    return os.path.join(ROOT, "oracle", "_ref", "ygz_ref_orb")


def run_reference_cpu(frames, threads, repeat, nfeatures=NFEAT):
    """Times the reference's own ORBextractor.cc (oracle/_ref/ygz_ref_orb) on host cores."""
    import tempfile
    with tempfile.TemporaryDirectory() as td:
        fin = os.path.join(td, "in.u8")
        frames.tofile(fin)
        cmd = [ref_binary(), "--in", fin, "--w", str(frames.shape[2]), "--h", str(frames.shape[1]), "--n", str(len(frames)),
               "--nfeatures", str(nfeatures),
               "--scale", repr(SCALE), "--levels", str(LEVELS), "--ini", str(INI_TH), "--min", str(MIN_TH),
               "--repeat", str(repeat), "--threads", str(threads)]
        out = subprocess.run(cmd, capture_output=True, text=True, check=True).stdout
    return json.loads(out.strip().splitlines()[-1])


def run_port_cpu(frames, threads, repeat):
    """Fallback CPU baseline: the oracle port (oracle/libygz_oracle.so), threads over frames."""
    import oracle_lib as O
    p = O.params(NFEAT, SCALE, LEVELS, INI_TH, MIN_TH)
    t0 = time.perf_counter()
    for _ in range(repeat):
        O.extract_batch(p, frames, threads=threads, want_output=False)
    sec = time.perf_counter() - t0
    return {"frames": len(frames) * repeat, "seconds": sec, "frames_per_s": len(frames) * repeat / sec, "threads": threads}


KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
SBP_TH, SBP_RATIO = 1.0, 0.8     # Tracking::SearchLocalPoints: ORBmatcher matcher(0.8); th = 1 (src/Tracking.cc:1662-1674)


def make_mappoints(counts, kps, desc, seed):
    """per frame N_MAPPOINTS synthetic MapPoints = random keypoints of the frame's
    extraction, projected with +-3 px jitter, viewCos ~ U(0.9, 1), level = source octave, descriptor =
    source descriptor with 0..40 random bit flips.  Arrays at fixed stride (F, N_MAPPOINTS)."""
    F, M = len(counts), N_MAPPOINTS
    rng = np.random.default_rng(seed)
    mp = {"proj_x": np.zeros((F, M), np.float32), "proj_y": np.zeros((F, M), np.float32),
          "proj_xr": np.zeros((F, M), np.float32), "view_cos": rng.uniform(0.9, 1.0, (F, M)).astype(np.float32),
          "level": np.zeros((F, M), np.int32), "in_view": np.ones((F, M), np.uint8), "has_obs": np.ones((F, M), np.uint8),
          "desc": np.zeros((F, M, 32), np.uint8)}
    for f in range(F):
        n = int(counts[f])
        if n == 0:
            mp["in_view"][f] = 0
            continue
        src = rng.integers(0, n, size=M)
        mp["proj_x"][f] = kps["x"][f, src] + rng.uniform(-3, 3, M).astype(np.float32)
        mp["proj_y"][f] = kps["y"][f, src] + rng.uniform(-3, 3, M).astype(np.float32)
        mp["level"][f] = kps["octave"][f, src]
        pflip = rng.integers(0, 41, size=M).astype(np.float32) / 256.0
        bits = rng.random((M, 256), dtype=np.float32) < pflip[:, None]
        mp["desc"][f] = desc[f, src] ^ np.packbits(bits, axis=1, bitorder="little")
    mp["n_mp"] = np.full(F, M, np.int32)
    return mp


def cpu_search_by_projection(counts, kps, desc, mp, threads, repeat):
    """Times ORBmatcher::SearchByProjection (threads over frames) on host cores: the reference's own src/ORBmatcher.cc +
    Frame::AssignFeaturesToGrid / GetFeaturesInArea (oracle/_ref/libygz_ref_match.so, compiled from the reference's source) when
    that library travelled with the snapshot, else the oracle port."""
    import oracle_lib as O
    import ref_match_lib as R
    import synth
    sf, _, _ = synth.scale_tables(SCALE, LEVELS)
    bounds = (0.0, 0.0, float(W), float(H))
    if R.available():
        _, nm, sec = R.search_by_projection_batch(counts, mp["n_mp"], kps, desc, bounds, sf, mp, SBP_TH, False, SBP_RATIO,
                                                  threads=threads, repeat=repeat)
        kind = "reference"
    else:
        t0 = time.perf_counter()
        for _ in range(repeat):
            _, nm = O.search_by_projection_batch(counts, mp["n_mp"], kps, desc, bounds, sf, mp, SBP_TH, False, SBP_RATIO, threads=threads)
        sec = time.perf_counter() - t0
        kind = "oracle-port"
    return {"frames": len(counts) * repeat, "seconds": sec, "frames_per_s": len(counts) * repeat / sec,
            "matches_per_frame": float(nm.mean()), "kind": kind}


def cpu_extract(frames, threads, repeat):
    if os.path.exists(ref_binary()):
        return "reference", run_reference_cpu(frames, threads, repeat)
    return "port", run_port_cpu(frames, threads, repeat)


def workload_config(B):
    """The `config` object of the JSON line: identical for both arms (the reference arm runs a bounded sample of it per step)."""
    M = N_MAPPOINTS
    return {"workload": "EuRoC-shape 752x480 mono frame stream, 1000 features + SearchByProjection against %d projected "
                        "MapPoints per frame (configs[1]); %d frames per step per GPU" % (M, B),
            "nfeatures": NFEAT, "scale": SCALE, "levels": LEVELS, "fast": [INI_TH, MIN_TH], "mappoints": M, "batch": B,
            "sbp": {"th": SBP_TH, "nnratio": SBP_RATIO},
            "l2": "inputs larger than L2 (%.0f MB of frames + %.0f MB of pyramid/blur workspace per step)"
                  % (B * W * H / 1e6, B * 2.0)}


def reference_sample_frames(threads):
    """Frames per CPU step: >= 8 per thread so that thread start-up and the tail of the dynamic schedule are amortised."""
    return max(8 * threads, 64)


def run_reference_arm(args, rank, world):
    """--impl reference: the reference's CPU implementation of the step on the box's host cores, rank 0 only:
    extraction = ORBextractor.cc itself (oracle/_ref/ygz_ref_orb, the reference's own source compiled unmodified), ONE process
    for all timed steps (--repeat steps); SearchByProjection = the reference's own ORBmatcher.cc (oracle/_ref/ygz_ref_match)
    when that binary exists, else the oracle port."""
    if rank != 0:
        return
    import oracle_lib as O
    threads = os.cpu_count() or 1
    sample = reference_sample_frames(threads)
    frames = make_frames(sample, 0)
    p = O.params(NFEAT, SCALE, LEVELS, INI_TH, MIN_TH)
    cap = NFEAT + 64
    counts, kps, desc = O.extract_batch(p, frames, threads=threads, cap=cap)    # untimed set-up: inputs of the search
    mp = make_mappoints(counts, kps, desc, 1)
    t0 = time.perf_counter()
    if args.warmup > 0:
        cpu_extract(frames, threads, args.warmup)
        cpu_search_by_projection(counts, kps, desc, mp, threads, 1)
    kind, st = cpu_extract(frames, threads, args.steps)                          # one process, `steps` passes over the sample
    sb = cpu_search_by_projection(counts, kps, desc, mp, threads, args.steps)
    wall = time.perf_counter() - t0
    t_ext, t_sbp = st["seconds"], sb["seconds"]
    fps = sample * args.steps / (t_ext + t_sbp)
    line = {
        "impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * (t_ext + t_sbp) / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(args.batch),
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": threads, "kind": kind,
                         "kind_per_leg": {"extraction": kind, "search_by_projection": sb["kind"]},
                         "sample": "each step = a bounded sample of %d frames of the workload (%d per thread), %d steps in one "
                                   "process, %d threads over frames; extraction = reference ORBextractor.cc (timed inside the "
                                   "binary), SearchByProjection = %s; wall %.1f s"
                                   % (sample, sample // threads, args.steps, threads, sb["kind"], wall)},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "hamming": {"mappoints_per_s": sample * args.steps * N_MAPPOINTS / t_sbp, "extract_frames_per_s": sample * args.steps / t_ext},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def sharded_block(args, Y, torch, dist, rank, local_rank, world, dev):
    """BASELINE.json configs[3] and [4], the two workloads north_star shards: through the library's multi-GPU layer
    (ygzorb_extract_batch_sharded / ygzorb_hamming_knn2_sharded: contiguous shards, the gather of the results on rank 0 over NCCL
    INSIDE the timed region).  Strong scaling: the totals are fixed, every N (including 1) runs the same job."""
    import synth
    comm = Y.Comm.from_torch_distributed(local_rank) if world > 1 else Y.Comm(0, 1, local_rank)

    def max_ms(ms):
        if dist is None:
            return ms
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
    out = {}
    # ---- configs[3]: 4096 distinct 1280x720 frames, 2000 features, streamed from pinned host memory, frame-sharded
    w, h, nf, n_total, chunk = 1280, 720, 2000, args.shard_frames, 256
    b, e = Y.shard_range(n_total, rank, world)
    n = e - b
    host = torch.empty((max(n, 1), h, w), dtype=torch.uint8).pin_memory()
    host.numpy()[:n] = synth.frame_batch(w, h, n, seed0=1000 + 16 * rank, distinct=16) if n else 0   # every frame differs (rolled textures)
    ex = Y.Extractor(nf, SCALE, LEVELS, INI_TH, MIN_TH, max_batch=chunk, device=local_rank)
    cap = ex.max_keypoints(w, h)
    stream = torch.cuda.ExternalStream(ex.stream(), device=dev)
    root = rank == 0
    g_counts = torch.zeros(n_total if root else 1, dtype=torch.int32, device=dev)
    g_off = torch.zeros(n_total if root else 1, dtype=torch.int64, device=dev)
    g_kps = torch.zeros(((n_total * cap) if root else 1, 7), dtype=torch.float32, device=dev)
    g_desc = torch.zeros(((n_total * cap) if root else 1, 32), dtype=torch.uint8, device=dev)
    l_counts = torch.zeros(max(n, 1), dtype=torch.int32, device=dev)
    l_kps = torch.zeros((max(n, 1) * cap, 7), dtype=torch.float32, device=dev)
    l_desc = torch.zeros((max(n, 1) * cap, 32), dtype=torch.uint8, device=dev)

    def run_sharded():
        comm.extract_batch_sharded(ex, host.data_ptr(), n_total, w, h, w, w * h, Y.NO_SYNC, cap, chunk, g_counts.data_ptr(),
                                   g_off.data_ptr(), g_kps.data_ptr(), g_desc.data_ptr())

    def run_local():
        if n:
            ex.extract_raw(host.data_ptr(), n, w, h, w, w * h, Y.OUT_DEVICE | Y.NO_SYNC, cap, l_counts.data_ptr(), l_kps.data_ptr(), l_desc.data_ptr())

    d_frames = host.to(dev)                                     # the same shard resident in HBM (3.8 GB / N): kernels + gather without the host link

    def run_sharded_dev():
        comm.extract_batch_sharded(ex, d_frames.data_ptr(), n_total, w, h, w, w * h, Y.IN_DEVICE | Y.NO_SYNC, cap, chunk, g_counts.data_ptr(),
                                   g_off.data_ptr(), g_kps.data_ptr(), g_desc.data_ptr())

    def timed(fn):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); fn(); e1.record(stream)
        ex.sync(); torch.cuda.synchronize()
        return max_ms(e0.elapsed_time(e1))
    run_sharded(); ex.sync()                                   # warm-up (NCCL connections, workspace growth)
    ms_local = timed(run_local)
    ms = min(timed(run_sharded), timed(run_sharded))
    run_sharded_dev(); ex.sync()
    ms_dev = min(timed(run_sharded_dev), timed(run_sharded_dev))
    n_kp_local = torch.tensor([int(l_counts[:n].sum().item())], dtype=torch.int64, device=dev)
    if dist is not None:
        dist.all_reduce(n_kp_local)
    if root:
        total_kp = int(g_counts.sum().item())
        assert total_kp == int(n_kp_local.item()) and total_kp > 1000 * n_total, "gathered counts differ from the shards' own"
        last = n_total - 1
        assert int(g_off[last].item()) >= 0
        out["config3"] = {
            "workload": "batched offline extraction of %d distinct 1280x720 frames, 2000 features each, from pinned host memory, "
                        "frame-sharded over %d GPU(s); gather of counts + packed records on rank 0 inside the timed region" % (n_total, world),
            "ms": ms, "frames_per_s": n_total / (ms * 1e-3), "ms_extract_only": ms_local, "gather_ms_exposed": max(ms - ms_local, 0.0),
            "h2d_GBps_all_ranks": n_total * w * h / (ms * 1e-3) / 1e9,
            "ms_device_resident": ms_dev, "frames_per_s_device_resident": n_total / (ms_dev * 1e-3),
            "gathered_bytes": total_kp * 60 + 4 * n_total, "keypoints_per_frame": total_kp / n_total, "chunk_frames": chunk, "scaling": "strong"}
        if world == 1 and not args.no_cpu and os.path.exists(ref_binary()):
            # the reference's own ORBextractor.cc on a bounded sample of the same frames, all host threads (a few seconds)
            threads = os.cpu_count() or 1
            ns = min(n, max(4 * threads, 32))
            r = run_reference_cpu(host.numpy()[:ns], threads, 4, nfeatures=nf)
            out["config3"]["cpu_baseline"] = {"value": r["frames_per_s"], "unit": "frames/s", "cores": threads, "kind": "reference",
                                              "sample": "%d of the 1280x720 frames x 4 passes = reference ORBextractor.cc (%.1f s)" % (ns, r["seconds"])}
    del host, d_frames, g_kps, g_desc, l_kps, l_desc
    ex.close()
    # ---- configs[4]: brute-force Hamming kNN-2 + ratio 0.6, nq x nt 256-bit rows, query-sharded; train rows broadcast from rank 0
    nq, nt = args.shard_queries, args.shard_queries
    b, e = Y.shard_range(nq, rank, world)
    n = e - b
    rng = np.random.default_rng(5)
    train = torch.from_numpy(rng.integers(0, 256, size=(nt, 32), dtype=np.uint8)).to(dev) if root else torch.zeros((nt, 32), dtype=torch.uint8, device=dev)
    ref_train = train
    if dist is not None:
        ref_train = train.clone()
        dist.broadcast(ref_train, src=0)                        # set-up only: the queries are derived from train rows
    g = torch.Generator(device=dev); g.manual_seed(100 + rank)
    query = torch.randint(0, 256, (max(n, 1), 32), dtype=torch.uint8, device=dev, generator=g)
    half = n // 2
    if half:
        src = torch.randint(0, nt, (half,), device=dev, generator=g)
        noise = (torch.rand((half, 32, 8), device=dev, generator=g) < 0.1).to(torch.uint8)
        bits = (noise << torch.arange(8, device=dev, dtype=torch.uint8)).sum(dim=2).to(torch.uint8)
        query[:half] = ref_train[src] ^ bits
    del ref_train
    mt = Y.Matcher(local_rank)
    mstream = torch.cuda.ExternalStream(mt.stream(), device=dev)
    res = [torch.full((nq,), -7, dtype=torch.int32, device=dev) for _ in range(4)]

    def run_knn():
        comm.hamming_knn2_sharded(mt, query.data_ptr(), nq, train.data_ptr(), nt, 0.6, 50, Y.BROADCAST_TRAIN | Y.NO_SYNC,
                                  *[r.data_ptr() for r in res])
    comm.hamming_knn2_sharded(mt, query.data_ptr(), nq, train.data_ptr(), nt if n else nt, 0.6, 50, Y.BROADCAST_TRAIN | Y.NO_SYNC,
                              *[r.data_ptr() for r in res])
    mt.sync()
    peak = mt.popc_peak()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(mstream); run_knn(); e1.record(mstream)
    mt.sync(); torch.cuda.synchronize()
    ms = max_ms(e0.elapsed_time(e1))
    if root:
        matches = int((res[3] >= 0).sum().item())
        assert int((res[1] == -7).sum().item()) == 0 and matches > nq // 4, "kNN gather incomplete"
        pairs = nq * nt / (ms * 1e-3)
        out["config4"] = {
            "workload": "brute-force Hamming kNN k=2 + ratio 0.6, %d x %d 256-bit rows, query-sharded over %d GPU(s); train rows "
                        "broadcast from rank 0 and (idx, d1, d2, match) gathered on rank 0 inside the timed region" % (nq, nt, world),
            "ms": ms, "pairs_per_s": pairs, "queries_per_s": nq / (ms * 1e-3), "matches": matches, "popc_peak_per_s_per_gpu": peak,
            "frac_of_popc_peak": pairs * 8 / (peak * world), "scaling": "strong"}
        if world == 1 and not args.no_cpu:
            # CPU figure beside it: the oracle port of the same rule (DescriptorDistance's bit trick per pair, top-2 + ratio),
            # threads over queries, on a bounded sample (sq x st rows) of the same arrays
            import oracle_lib as O
            threads = os.cpu_count() or 1
            sq, st = min(nq, 4096 * threads), min(nt, 65536)
            qh, th_ = query[:sq].cpu().numpy(), train[:st].cpu().numpy()
            t0 = time.perf_counter()
            O.hamming_knn2(qh, th_, 0.6, 50, threads=threads)
            sec = time.perf_counter() - t0
            out["config4"]["cpu_baseline"] = {"value": sq * st / sec, "unit": "pairs/s", "cores": threads, "kind": "port",
                                              "sample": "%d x %d rows of the same arrays, oracle port (%.1f s)" % (sq, st, sec)}
    mt.close()
    comm.close()
    return out


def run_ours(args, rank, local_rank, world):
    import torch
    import synth
    import ygzorb_ctypes as Y
    if not torch.cuda.is_available() or Y.device_count() < 1:
        raise RuntimeError("bench.py needs a CUDA device: the product has no CPU path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa = bind_to_gpu_cpus(local_rank) if world > 1 else None
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    B, M = args.batch, N_MAPPOINTS
    frames_np = make_frames(B, seed0=100 * rank)
    host_in = torch.empty((B, H, W), dtype=torch.uint8).pin_memory()
    host_in.numpy()[:] = frames_np
    dev_in = host_in.to(dev)
    ex = Y.Extractor(NFEAT, SCALE, LEVELS, INI_TH, MIN_TH, max_batch=B, device=local_rank)
    # host-input calls of the e2e leg: three upload chunks per step (171 frames of 512) measured best (256: -3 %, 128: -3 %, 64: -8 %):
    # a MapPoint upload waits behind at most one 1.1 ms frame chunk, and the kernels still run on batches large enough
    ex.set_pipeline(args.pipe if args.pipe >= 0 else (args.batch + 2) // 3)
    mt = Y.Matcher(local_rank)
    mt.set_stream(ex.stream())          # the search consumes the extraction's device-resident outputs in stream order
    cap = ex.max_keypoints(W, H)
    d_counts = torch.zeros(B, dtype=torch.int32, device=dev)
    d_kps = torch.zeros((B, cap, 7), dtype=torch.float32, device=dev)
    d_desc = torch.zeros((B, cap, 32), dtype=torch.uint8, device=dev)
    d_kp2mp = torch.zeros((B, cap), dtype=torch.int32, device=dev)
    d_best = torch.zeros((B, M), dtype=torch.int32, device=dev)
    d_nm = torch.zeros(B, dtype=torch.int32, device=dev)
    stream = torch.cuda.ExternalStream(ex.stream(), device=dev)
    # ---- set-up (untimed): one extraction to derive the synthetic MapPoints of every frame
    ex.extract_raw(dev_in.data_ptr(), B, W, H, W, W * H, Y.IN_DEVICE | Y.OUT_DEVICE, cap,
                   d_counts.data_ptr(), d_kps.data_ptr(), d_desc.data_ptr())
    counts0 = d_counts.cpu().numpy()
    kps0 = d_kps.cpu().numpy().view(KP_DTYPE).reshape(B, cap)
    desc0 = d_desc.cpu().numpy()
    mp = make_mappoints(counts0, kps0, desc0, 1 + rank)
    # the frames are monocular (mvuRight = -1, u_right = NULL below), so mTrackProjXR is never read (src/ORBmatcher.cc:93-97) and is
    # neither marshalled nor uploaded: proj_xr = NULL in ygzorb_mappoint_proj
    mp_fields = ["n_mp", "proj_x", "proj_y", "proj_xr", "view_cos", "level", "in_view", "has_obs", "desc"]
    mp_keys = [k for k in mp_fields if k != "proj_xr"]
    h_mp = {k: torch.from_numpy(mp[k]).pin_memory() for k in mp_keys}
    d_mp = {k: h_mp[k].to(dev) for k in mp_keys}

    def mp_struct(m):
        return Y.MapPointProj(M, *[m[k].data_ptr() if k in m else None for k in mp_fields])
    sf, _, _ = synth.scale_tables(SCALE, LEVELS)
    assert sf.tobytes() == ex.tables()["scale"].tobytes()
    F = Y.FrameFeatures(B, cap, d_counts.data_ptr(), d_kps.data_ptr(), d_desc.data_ptr(), None, None,
                        0.0, 0.0, float(W), float(H), LEVELS, sf.ctypes.data)
    P = mp_struct(d_mp)
    dflags = Y.IN_DEVICE | Y.OUT_DEVICE | Y.NO_SYNC
    d_out = {"counts": d_counts, "kps": d_kps, "desc": d_desc, "kp2mp": d_kp2mp, "nm": d_nm}

    def search(F=F, P=P, out=d_out, best=d_best):
        mt.search_by_projection_raw(F, P, SBP_TH, False, SBP_RATIO, dflags, out["kp2mp"].data_ptr(), best.data_ptr(),
                                    out["nm"].data_ptr())

    def step_device():
        ex.extract_raw(dev_in.data_ptr(), B, W, H, W, W * H, dflags, cap, d_counts.data_ptr(), d_kps.data_ptr(), d_desc.data_ptr())
        search()

    # ---- end-to-end leg: frames and MapPoints come from pinned host memory, every result goes back to pinned
    # host memory, each step.  Two slots of device buffers so that consecutive steps pipeline: the library
    # uploads the frames in chunks on its own stream beside the kernels (ygzorb_extractor_set_pipeline); the
    # MapPoint upload and the result download run on two side streams here, ordered by events.
    copy_in, copy_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    slots = []
    N_SLOTS = args.slots
    for sl in range(N_SLOTS):
        o = d_out if sl == 0 else {k: torch.zeros_like(v) for k, v in d_out.items()}
        m = d_mp if sl == 0 else {k: torch.zeros_like(v) for k, v in d_mp.items()}
        slots.append({
            "out": o, "mp": m, "best": d_best if sl == 0 else torch.zeros_like(d_best),
            "F": Y.FrameFeatures(B, cap, o["counts"].data_ptr(), o["kps"].data_ptr(), o["desc"].data_ptr(), None, None,
                                 0.0, 0.0, float(W), float(H), LEVELS, sf.ctypes.data),
            "P": mp_struct(m),
            "h_out": {k: torch.zeros(v.shape, dtype=v.dtype).pin_memory() for k, v in d_out.items()},
            "mp_ready": torch.cuda.Event(), "search_done": torch.cuda.Event(), "out_done": torch.cuda.Event()})
    h_out = slots[0]["h_out"]
    host_step = [0]
    timeline = [] if args.timeline else None   # --timeline: per-step event times of the e2e leg (diagnostic, on stderr)

    up_keys = [list(mp_keys)]                           # the MapPoint fields uploaded every step (all of them: the headline e2e)

    def step_host():
        S = slots[host_step[0] % N_SLOTS]
        host_step[0] += 1
        copy_in.wait_event(S["search_done"])            # the previous search on this slot has read its MapPoints
        # The upload is NOT held back until the previous step's search is done (round 1 did that): the copy engine serves copies in
        # the order they become ready, and a MapPoint upload that becomes ready late queues behind up to three frame chunks (5 ms),
        # stalling its search and, behind it in the stream, the next extraction (bench.py --timeline shows the bubbles).
        tl = [torch.cuda.Event(enable_timing=True) for _ in range(5)] if timeline is not None else None
        with torch.cuda.stream(copy_in):
            if tl: tl[0].record(copy_in)
            for k in up_keys[0]:
                S["mp"][k].copy_(h_mp[k], non_blocking=True)
            S["mp_ready"].record(copy_in)
            if tl: tl[1].record(copy_in)
        stream.wait_event(S["out_done"])                # the previous results of this slot are on the host
        o = S["out"]
        ex.extract_raw(host_in.data_ptr(), B, W, H, W, W * H, Y.OUT_DEVICE | Y.NO_SYNC, cap,
                       o["counts"].data_ptr(), o["kps"].data_ptr(), o["desc"].data_ptr())
        if tl: tl[2].record(stream)
        stream.wait_event(S["mp_ready"])
        search(S["F"], S["P"], o, S["best"])
        S["search_done"].record(stream)
        if tl: tl[3].record(stream)
        copy_out.wait_event(S["search_done"])
        with torch.cuda.stream(copy_out):
            for k in o:
                S["h_out"][k].copy_(o[k], non_blocking=True)
            S["out_done"].record(copy_out)
            if tl: tl[4].record(copy_out)
        if tl: timeline.append(tl)

    def drain_host():                                    # the timed region ends when the last download has landed
        for S in slots:
            stream.wait_event(S["out_done"])

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        ex.sync()

    def timed(step, steps, drain=None):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = ex.launch_count() + mt.launch_count()
        e0.record(stream)
        for _ in range(steps):
            step()
        if drain is not None:
            drain()
        e1.record(stream)
        ex.sync()
        torch.cuda.synchronize()
        barrier()
        ms = e0.elapsed_time(e1)
        if dist is not None:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, ex.launch_count() + mt.launch_count() - l0

    for _ in range(args.warmup):
        step_device()
    ex.sync()
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms, launches = timed(step_device, args.steps)
    clocks = sampler.stop()
    n_kp = int(d_counts.sum().item())
    n_match = int(d_nm.sum().item())
    for _ in range(max(args.warmup, 2)):
        step_host()
    ex.sync()
    torch.cuda.synchronize()
    if timeline is not None:
        timeline.clear()
    ms_e2e, _ = timed(step_host, args.steps, drain_host)
    if timeline:
        t0 = timeline[0][0]
        sys.stderr.write("e2e timeline, ms since the first MapPoint upload began: [mp_up start, mp_up end, extraction end, search end, download end]\n")
        for tl in timeline:
            sys.stderr.write("  " + " ".join("%7.3f" % t0.elapsed_time(e) for e in tl) + "\n")
        timeline = None
    for S in slots:
        assert int(S["h_out"]["counts"].sum().item()) == n_kp and int(S["h_out"]["nm"].sum().item()) == n_match, \
            "host and device paths disagree"
    # variant: the local map's descriptors stay resident in HBM between frames (a MapPoint's descriptor changes only when
    # ComputeDistinctiveDescriptors reruns); per step only the projections go up (22 of 54 B per MapPoint).  Reported beside the
    # headline e2e, which re-sends everything.
    up_keys[0] = [k for k in mp_keys if k != "desc"]
    for _ in range(2):
        step_host()
    ex.sync(); torch.cuda.synchronize()
    ms_e2e_res, _ = timed(step_host, args.steps, drain_host)
    for S in slots:
        assert int(S["h_out"]["nm"].sum().item()) == n_match, "resident-map variant disagrees"
    up_keys[0] = list(mp_keys)
    # the floor of the e2e leg: this step's uploads alone (frames + MapPoint records from pinned memory into HBM), nothing else
    # running, all ranks at once — what the host side of the box can deliver
    up_bytes = B * W * H + sum(int(h_mp[k].numel() * h_mp[k].element_size()) for k in mp_keys)

    def step_upload_only():
        with torch.cuda.stream(copy_in):
            dev_in.copy_(host_in, non_blocking=True)
            for k in mp_keys:
                d_mp[k].copy_(h_mp[k], non_blocking=True)
            slots[0]["mp_ready"].record(copy_in)
        stream.wait_event(slots[0]["mp_ready"])
    for _ in range(2):
        step_upload_only()
    torch.cuda.synchronize()
    ms_floor, _ = timed(step_upload_only, args.steps)
    e2e_extra = {"value_resident_map": world * B * args.steps / (ms_e2e_res * 1e-3),
                 "resident_map_note": "same step with the MapPoint descriptors resident in HBM; per step the frames and 22 B of projection "
                                      "data per MapPoint go up",
                 "h2d_bytes_per_step_resident_map": B * W * H + sum(int(h_mp[k].numel() * h_mp[k].element_size()) for k in mp_keys if k != "desc"),
                 "h2d_floor_ms_per_step": ms_floor / args.steps,
                 "h2d_floor_GBps_all_ranks": world * up_bytes * args.steps / (ms_floor * 1e-3) / 1e9,
                 "h2d_achieved_GBps_all_ranks": world * up_bytes * args.steps / (ms_e2e * 1e-3) / 1e9,
                 "h2d_floor_frac": ms_floor / ms_e2e,
                 "h2d_floor_note": "floor = the step's uploads alone from pinned memory on all ranks at once (max over ranks); "
                                   "h2d_floor_frac = floor time / e2e time: 1.0 means the e2e leg runs at the speed of its uploads"}
    # per-stage device times: CUDA events between the extractor's stages + around the search, on the shared stream
    ex.set_profiling(True)
    acc = np.zeros(8)
    reps = 3
    for _ in range(reps):
        ex.extract_raw(dev_in.data_ptr(), B, W, H, W, W * H, dflags, cap, d_counts.data_ptr(), d_kps.data_ptr(), d_desc.data_ptr())
        ex.sync()
        acc += ex.stage_ms()
    ex.set_profiling(False)
    stage_ms = {STAGE_NAMES[i]: float(acc[i] / reps) for i in range(7)}
    ms_sbp, _ = timed(search, args.steps)
    stage_ms["search_by_projection"] = ms_sbp / args.steps
    # ---- second workload, same shapes: frames with low-contrast halves and a flat band (tests/synth.mixed_frame), so that a
    # large share of the FAST cells takes the minThFAST retry of src/ORBextractor.cc:764-770 (the dense texture above never does);
    # the buffers of the first workload are refilled in place and the same two timed loops run again
    mixed = None
    if not args.no_mixed:
        mixed_np = make_mixed_frames(B, 500 + 100 * rank)
        host_in.numpy()[:] = mixed_np
        dev_in.copy_(host_in)
        torch.cuda.synchronize()
        ex.extract_raw(dev_in.data_ptr(), B, W, H, W, W * H, Y.IN_DEVICE | Y.OUT_DEVICE, cap,
                       d_counts.data_ptr(), d_kps.data_ptr(), d_desc.data_ptr())
        counts_m = d_counts.cpu().numpy()
        kps_m = d_kps.cpu().numpy().view(KP_DTYPE).reshape(B, cap)
        desc_m = d_desc.cpu().numpy()
        mp_m = make_mappoints(counts_m, kps_m, desc_m, 7 + rank)
        for k in mp_keys:
            h_mp[k].numpy()[...] = mp_m[k]
            d_mp[k].copy_(h_mp[k])
        torch.cuda.synchronize()
        for _ in range(3):
            step_device()
        ex.sync()
        ms_mixed, _ = timed(step_device, args.steps)
        n_kp_m, n_match_m = int(d_counts.sum().item()), int(d_nm.sum().item())
        for _ in range(2):
            step_host()
        ex.sync(); torch.cuda.synchronize()
        ms_e2e_mixed, _ = timed(step_host, args.steps, drain_host)
        for S in slots:
            assert int(S["h_out"]["counts"].sum().item()) == n_kp_m and int(S["h_out"]["nm"].sum().item()) == n_match_m, \
                "host and device paths disagree on the mixed workload"
        mixed = {"workload": "same shapes on frames whose right half has 0.18 of the contrast and which carry a flat band",
                 "value": world * B * args.steps / (ms_mixed * 1e-3), "unit": "frames/s", "ms_per_step": ms_mixed / args.steps,
                 "e2e": world * B * args.steps / (ms_e2e_mixed * 1e-3), "e2e_ms_per_step": ms_e2e_mixed / args.steps,
                 "keypoints_per_frame": n_kp_m / B, "mappoint_matches_per_frame": n_match_m / B}
        if rank == 0:
            mixed.update(retry_cell_fraction(mixed_np[0]))
            if world == 1 and not args.no_cpu:
                mixed["cpu_baseline"] = cpu_baseline_leg(mixed_np, counts_m, kps_m, desc_m, mp_m, budget_s=5.0)
    def shutdown():
        """Orderly teardown, normal interpreter exit afterwards (exit hooks run, the mapped libraries stay visible):
        everything queued on the handles' streams is drained, torch's work is drained, then the handles go — BEFORE any
        pinned buffer is released, which is the order an embedder gets when the handle is a member destroyed first."""
        ex.sync()
        mt.sync()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        mt.close()
        ex.close()
        torch.cuda.synchronize()

    sharded = None
    if not args.no_sharded:
        # free the headline workload's device buffers first: configs[3] holds ~1.5 GB of results on rank 0
        for S in slots:
            S.clear()
        del slots[:]
        sharded = sharded_block(args, Y, torch, dist, rank, local_rank, world, dev)
    if rank != 0:
        shutdown()
        return
    line = build_line(args, world, B, cap, ms, ms_e2e, launches, clocks, stage_ms, n_kp, n_match, h_mp, mp_keys)
    line["e2e"].update(e2e_extra)
    if numa is not None:
        line["config"]["rank_cpu_binding"] = numa
    if world == 1 and not args.no_cpu:
        line["hamming"].update(knn_probe(Y, mt, torch))
        line["cpu_baseline"] = cpu_baseline_leg(frames_np, counts0, kps0, desc0, mp)
    if mixed is not None:
        line["mixed"] = mixed
    if sharded is not None:
        line["sharded"] = sharded
    line["native_library"] = Y.LIB_PATH if os.path.exists(Y.LIB_PATH) else None
    print(json.dumps(line), flush=True)
    shutdown()


def build_line(args, world, B, cap, ms, ms_e2e, launches, clocks, stage_ms, n_kp, n_match, h_mp, mp_keys):
    hbm_peak, peak_src = peaks()
    M = N_MAPPOINTS
    fps = world * B * args.steps / (ms * 1e-3)
    fps_e2e = world * B * args.steps / (ms_e2e * 1e-3)
    # dominant kernel = the slowest extraction stage; algorithmic bytes per launch from SURVEY.md §8(d)
    kern_stages = {"pyramid": "pyramid", "fast": "fast", "blur": "blur", "quadtree": "quadtree_orient_desc",
                   "orient_desc": "quadtree_orient_desc"}
    dom = max(kern_stages, key=lambda k: stage_ms[k])
    launches_per_stage = {"pyramid": LEVELS - 1}.get(dom, 1)
    dom_bytes = STAGE_BYTES[kern_stages[dom]] * B
    achieved = dom_bytes / (stage_ms[dom] * 1e-3) / 1e9
    staged = STAGED_TOTAL * B * args.steps / (ms * 1e-3) / 1e9
    # DRAM traffic of that kernel per launch: dram__bytes_read + dram__bytes_write of one `ncu --set full` capture
    # (tools/ncu_traffic.py -> profiles/r2_traffic.json, per frame), scaled to this launch's frames
    traffic, traffic_src = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json")))
        tk = tj["kernels"]["k_" + dom]
        traffic = (tk["dram_read_bytes_per_frame"] + tk["dram_write_bytes_per_frame"]) * B / launches_per_stage
        traffic_src = "profiles/r2_traffic.json (%s, %d frames, scaled per frame)" % (tj["report"], tj["frames"])
    except (OSError, KeyError, ValueError):
        pass
    roofline = {"bound": "hbm", "kernel": "k_" + dom, "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                "frac": achieved / hbm_peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "note": "the kernel is ALU-pipe / issue bound (ncu: 85 % issue-slot utilisation, DRAM 4 %), see DESIGN.md 5",
                "launches_in_stage": launches_per_stage, "algorithmic_bytes_per_launch": dom_bytes / launches_per_stage,
                "pipeline_staged_GBps": staged, "pipeline_frac": staged / hbm_peak,
                # the streaming stages beside the dominant one: algorithmic bytes of the stage / its event time
                "per_stage": {k: {"achieved": STAGE_BYTES[kern_stages[k]] * B / (stage_ms[k] * 1e-3) / 1e9,
                                  "frac": STAGE_BYTES[kern_stages[k]] * B / (stage_ms[k] * 1e-3) / 1e9 / hbm_peak}
                              for k in ("pyramid", "fast", "blur")}}
    mp_bytes = sum(int(h_mp[k].numel() * h_mp[k].element_size()) for k in mp_keys)
    return {
        "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": workload_config(B),
        "clocks": clocks,
        "e2e": {"value": fps_e2e, "unit": "frames/s", "h2d_bytes_per_step": B * W * H + mp_bytes,
                "d2h_bytes_per_step": B * (4 + cap * 60 + cap * 4 + 4), "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": launches,
        "roofline": roofline,
        "stage_ms_per_step": stage_ms,
        "keypoints_per_frame": n_kp / B,
        "hamming": {"mappoints_per_s": world * B * M / (stage_ms["search_by_projection"] * 1e-3),
                    "mappoint_matches_per_frame": n_match / B, "sbp_ms_per_step": stage_ms["search_by_projection"]},
    }


def knn_probe(Y, mt, torch):
    """Bounded sample of BASELINE.json config 5 (brute-force kNN-2 + ratio): 32768 queries x 1M train rows on one GPU,
    against the measured POPC peak (8 POPC32 per 256-bit pair)."""
    import synth
    nq, nt = 32768, 1 << 20
    q, t = synth.knn_case(nq, nt, 1)
    dq, dt = torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda()
    out = [torch.zeros(nq, dtype=torch.int32, device="cuda") for _ in range(4)]
    mt.sync()
    peak = mt.popc_peak()
    stream = torch.cuda.ExternalStream(mt.stream())
    flags = Y.IN_DEVICE | Y.OUT_DEVICE | Y.NO_SYNC
    for _ in range(2):
        mt.hamming_knn2_raw(dq.data_ptr(), nq, dt.data_ptr(), nt, 0.6, 50, flags, *[o.data_ptr() for o in out])
    mt.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    reps = 3
    for _ in range(reps):
        mt.hamming_knn2_raw(dq.data_ptr(), nq, dt.data_ptr(), nt, 0.6, 50, flags, *[o.data_ptr() for o in out])
    e1.record(stream)
    mt.sync()
    ms = e0.elapsed_time(e1) / reps
    pairs = nq * nt / (ms * 1e-3)
    return {"knn2": {"workload": "%d x %d 256-bit rows, k=2 + ratio 0.6 (bounded sample of configs[4])" % (nq, nt),
                     "pairs_per_s": pairs, "ms": ms, "popc_peak_per_s": peak, "bound": "int-pipe (POPC)",
                     "frac_of_popc_peak": pairs * 8 / peak, "matches": int((out[3] >= 0).sum().item())}}


def cpu_baseline_leg(frames_np, counts, kps, desc, mp, budget_s=12.0):
    """The reference's CPU path on the box's host cores, bounded sample of the same workload."""
    threads = os.cpu_count() or 1
    n = min(reference_sample_frames(threads), len(frames_np))
    sample = frames_np[:n]
    kind, st1 = cpu_extract(sample, threads, 1)
    repeat = int(max(1, min(200, budget_s / max(st1["seconds"], 1e-3))))
    kind, st = cpu_extract(sample, threads, repeat)
    sub = {k: np.ascontiguousarray(v[:n]) for k, v in mp.items()}
    sb1 = cpu_search_by_projection(counts[:n].copy(), np.ascontiguousarray(kps[:n]), np.ascontiguousarray(desc[:n]), sub, threads, 1)
    rep2 = int(max(1, min(200, 3.0 / max(sb1["seconds"], 1e-4))))
    sb = cpu_search_by_projection(counts[:n].copy(), np.ascontiguousarray(kps[:n]), np.ascontiguousarray(desc[:n]), sub, threads, rep2)
    fps = 1.0 / (1.0 / st["frames_per_s"] + 1.0 / sb["frames_per_s"])
    _, s1 = cpu_extract(sample[:8], 1, 1)
    return {"value": fps, "unit": "frames/s", "cores": threads, "kind": kind,
            "kind_per_leg": {"extraction": kind, "search_by_projection": sb["kind"]},
            "sample": "%d frames: extraction x %d repeats = reference ORBextractor.cc (%.1f s, %.1f frames/s), SearchByProjection "
                      "x %d repeats = %s (%.1f s, %.0f frames/s); threads over frames"
                      % (n, repeat, st["seconds"], st["frames_per_s"], rep2, sb["kind"], sb["seconds"], sb["frames_per_s"]),
            "extract_frames_per_s": st["frames_per_s"], "sbp_mappoints_per_s": sb["frames_per_s"] * N_MAPPOINTS,
            "single_thread_extract_frames_per_s": s1["frames_per_s"]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--slots", type=int, default=N_SLOTS, help="device buffer sets of the e2e leg")
    ap.add_argument("--timeline", action="store_true", help="print per-step event times of the e2e leg to stderr")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline and kNN legs")
    ap.add_argument("--no-mixed", action="store_true", help="skip the second (mixed-contrast) workload")
    ap.add_argument("--no-sharded", action="store_true", help="skip the sharded block (configs[3] and [4] through the multi-GPU layer)")
    ap.add_argument("--shard-frames", type=int, default=4096, help="frames of configs[3]")
    ap.add_argument("--shard-queries", type=int, default=1 << 20, help="query = train rows of configs[4]")
    ap.add_argument("--pipe", type=int, default=-1, help="frames per pipelined upload chunk of the e2e leg (-1: a third of the batch; 0: one chunk)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return
    run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
