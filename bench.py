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


def ref_binary():
    return os.path.join(ROOT, "oracle", "_ref", "ygz_ref_orb")


def run_reference_cpu(frames, threads, repeat):
    """Times the reference's own ORBextractor.cc (oracle/_ref/ygz_ref_orb) on host cores."""
    import tempfile
    with tempfile.TemporaryDirectory() as td:
        fin = os.path.join(td, "in.u8")
        frames.tofile(fin)
        cmd = [ref_binary(), "--in", fin, "--w", str(W), "--h", str(H), "--n", str(len(frames)), "--nfeatures", str(NFEAT),
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


def cpu_extract(frames, threads, repeat):
    if os.path.exists(ref_binary()):
        return "reference", run_reference_cpu(frames, threads, repeat)
    return "port", run_port_cpu(frames, threads, repeat)


def run_reference_arm(args, rank, world):
    """--impl reference: the reference's CPU implementation on the box's host cores, rank 0 only."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    sample = max(2 * threads, 16)
    frames = make_frames(sample, 0)
    kind, _ = cpu_extract(frames[: max(threads, 4)], threads, 1)   # warm-up of page cache / binary
    for _ in range(max(args.warmup - 1, 0)):
        cpu_extract(frames, threads, 1)
    t0 = time.perf_counter()
    inner = 0.0
    for _ in range(args.steps):
        _, st = cpu_extract(frames, threads, 1)
        inner += st["seconds"]
    wall = time.perf_counter() - t0
    fps = sample * args.steps / inner
    line = {
        "impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * inner / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "EuRoC-shape 752x480 mono frame stream, 1000 features (configs[1]); bounded sample of "
                               "%d frames per step" % sample, "nfeatures": NFEAT, "scale": SCALE, "levels": LEVELS,
                   "fast": [INI_TH, MIN_TH]},
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": threads, "kind": kind,
                         "sample": "%d frames x %d steps, %d threads over frames; timed inside the binary "
                                   "(file I/O excluded), wall %.1f s" % (sample, args.steps, threads, wall)},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_ours(args, rank, local_rank, world):
    import torch
    import ygzorb_ctypes as Y
    if not torch.cuda.is_available() or Y.device_count() < 1:
        raise RuntimeError("bench.py needs a CUDA device: the product has no CPU path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    B = args.batch
    frames_np = make_frames(B, seed0=100 * rank)
    host_in = torch.empty((B, H, W), dtype=torch.uint8).pin_memory()
    host_in.numpy()[:] = frames_np
    dev_in = host_in.to(dev)
    ex = Y.Extractor(NFEAT, SCALE, LEVELS, INI_TH, MIN_TH, max_batch=B, device=local_rank)
    cap = ex.max_keypoints(W, H)
    d_counts = torch.zeros(B, dtype=torch.int32, device=dev)
    d_kps = torch.zeros((B, cap, 7), dtype=torch.float32, device=dev)
    d_desc = torch.zeros((B, cap, 32), dtype=torch.uint8, device=dev)
    h_counts = torch.zeros(B, dtype=torch.int32).pin_memory()
    h_kps = torch.zeros((B, cap, 7), dtype=torch.float32).pin_memory()
    h_desc = torch.zeros((B, cap, 32), dtype=torch.uint8).pin_memory()
    stream = torch.cuda.ExternalStream(ex.stream(), device=dev)

    def step_device():
        ex.extract_raw(dev_in.data_ptr(), B, W, H, W, W * H, Y.IN_DEVICE | Y.OUT_DEVICE | Y.NO_SYNC, cap,
                       d_counts.data_ptr(), d_kps.data_ptr(), d_desc.data_ptr())

    def step_host():
        ex.extract_raw(host_in.data_ptr(), B, W, H, W, W * H, Y.NO_SYNC, cap,
                       h_counts.data_ptr(), h_kps.data_ptr(), h_desc.data_ptr())

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        ex.sync()

    def timed(step, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = ex.launch_count()
        e0.record(stream)
        for _ in range(steps):
            step()
        e1.record(stream)
        ex.sync()
        barrier()
        ms = e0.elapsed_time(e1)
        if dist is not None:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, ex.launch_count() - l0

    for _ in range(args.warmup):
        step_device()
    ex.sync()
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms, launches = timed(step_device, args.steps)
    clocks = sampler.stop()
    n_kp = int(d_counts.sum().item())
    # end to end through the C ABI with host buffers (H2D of the frames + D2H of the results every step)
    for _ in range(max(args.warmup // 2, 1)):
        step_host()
    ex.sync()
    ms_e2e, _ = timed(step_host, args.steps)
    assert int(h_counts.sum().item()) == n_kp, "host and device paths disagree"
    # per-stage device times (CUDA events between the stages, on the extractor's stream)
    ex.set_profiling(True)
    acc = np.zeros(8)
    reps = 3
    for _ in range(reps):
        step_device()
        ex.sync()
        acc += ex.stage_ms()
    ex.set_profiling(False)
    stage_ms = {STAGE_NAMES[i]: float(acc[i] / reps) for i in range(7)}
    if rank != 0:
        return
    hbm_peak, peak_src = peaks()
    fps = world * B * args.steps / (ms * 1e-3)
    fps_e2e = world * B * args.steps / (ms_e2e * 1e-3)
    # dominant kernel = the slowest stage; its algorithmic bytes per launch from SURVEY.md §8(d)
    kern_stages = {"pyramid": "pyramid", "fast": "fast", "blur": "blur", "quadtree": "quadtree_orient_desc",
                   "orient_desc": "quadtree_orient_desc"}
    dom = max(kern_stages, key=lambda k: stage_ms[k])
    launches_per_stage = {"pyramid": LEVELS - 1}.get(dom, 1)
    dom_bytes = STAGE_BYTES[kern_stages[dom]] * B
    achieved = dom_bytes / (stage_ms[dom] * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                "frac": achieved / hbm_peak, "traffic": None, "peak_source": peak_src,
                "launches_in_stage": launches_per_stage,
                "algorithmic_bytes_per_launch": dom_bytes / launches_per_stage,
                "pipeline_staged_GBps": STAGED_TOTAL * B * args.steps / (ms * 1e-3) / 1e9,
                "pipeline_frac": STAGED_TOTAL * B * args.steps / (ms * 1e-3) / 1e9 / hbm_peak}
    line = {
        "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {"workload": "EuRoC-shape 752x480 mono frame stream, 1000 features (configs[1]); %d frames per step "
                               "per GPU" % B, "nfeatures": NFEAT, "scale": SCALE, "levels": LEVELS, "fast": [INI_TH, MIN_TH],
                   "batch": B, "l2": "inputs larger than L2 (%.0f MB of frames + %.0f MB of pyramid/blur workspace per step)"
                                    % (B * W * H / 1e6, B * 2.0)},
        "clocks": clocks,
        "e2e": {"value": fps_e2e, "unit": "frames/s", "h2d_bytes_per_step": B * W * H,
                "d2h_bytes_per_step": B * (4 + cap * 60), "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": launches,
        "roofline": roofline,
        "stage_ms_per_step": stage_ms,
        "keypoints_per_frame": n_kp / B,
    }
    if world == 1 and not args.no_cpu:
        threads = os.cpu_count() or 1
        sample = frames_np[: max(2 * threads, 16)]
        kind, st1 = cpu_extract(sample, threads, 1)
        repeat = int(max(1, min(50, 12.0 / max(st1["seconds"], 1e-3))))
        kind, st = cpu_extract(sample, threads, repeat)
        line["cpu_baseline"] = {"value": st["frames_per_s"], "unit": "frames/s", "cores": threads, "kind": kind,
                                "sample": "%d frames x %d repeats of the same workload, threads over frames (%.1f s)"
                                          % (len(sample), repeat, st["seconds"])}
        kind1, s1 = cpu_extract(sample[:8], 1, 1)
        line["cpu_baseline"]["single_thread_frames_per_s"] = s1["frames_per_s"]
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
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
