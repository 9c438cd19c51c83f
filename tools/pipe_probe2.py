"""Library pipeline alone: host frames in (pinned), results device-resident, no search: frames/s for several chunk sizes."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "orb-ygz-slam_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, bench
import ygzorb_ctypes as Y
B = 512
dev = torch.device("cuda", 0)
frames = bench.make_frames(B, 0)
host_in = torch.from_numpy(frames).pin_memory()
dev_in = host_in.to(dev)
for pipe in (256, 512):
    ex = Y.Extractor(bench.NFEAT, bench.SCALE, bench.LEVELS, bench.INI_TH, bench.MIN_TH, max_batch=B)
    ex.set_pipeline(pipe)
    cap = ex.max_keypoints(bench.W, bench.H)
    outs = [(torch.zeros(B, dtype=torch.int32, device=dev), torch.zeros((B, cap, 7), device=dev), torch.zeros((B, cap, 32), dtype=torch.uint8, device=dev)) for _ in range(2)]
    stream = torch.cuda.ExternalStream(ex.stream(), device=dev)
    for mode, src, fl in (("host", host_in, Y.OUT_DEVICE | Y.NO_SYNC), ("device", dev_in, Y.IN_DEVICE | Y.OUT_DEVICE | Y.NO_SYNC)):
        def step(i):
            c, k, d = outs[i & 1]
            ex.extract_raw(src.data_ptr(), B, bench.W, bench.H, bench.W, bench.W * bench.H, fl, cap, c.data_ptr(), k.data_ptr(), d.data_ptr())
        for i in range(3):
            step(i)
        ex.sync(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        t0 = time.perf_counter()
        for i in range(10):
            step(i)
        t_enq = time.perf_counter() - t0
        e1.record(stream); ex.sync(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print("pipe %d %-6s: %.3f ms per step, %.1fk frames/s (host enqueue %.2f ms per step)" % (pipe, mode, ms, B / ms, t_enq * 100))
    ex.sync(); torch.cuda.synchronize()
torch.cuda.synchronize()
Y.close_all()   # orderly shutdown: handles first, then a normal interpreter exit (exit hooks run)
