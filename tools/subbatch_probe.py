"""How many frames per kernel launch?  512 device-resident 752x480 frames through ygzorb_extract_batch with the extractor's
max_batch (= frames per launch of every stage) swept: small sub-batches keep pyramid + blurred pyramid (2.0 MB per frame) in
the 126 MB L2 between the stages, large ones amortise launches."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb-ygz-slam_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import synth  # noqa: E402
import ygzorb_ctypes as Y  # noqa: E402

W, H, B = 752, 480, 512
frames = torch.from_numpy(synth.frame_batch(W, H, B, seed0=0, distinct=8)).cuda()
out = {}
for mb in [512, 256, 128, 64, 48, 32, 24, 16, 8]:
    ex = Y.Extractor(1000, max_batch=mb)
    cap = ex.max_keypoints(W, H)
    counts = torch.zeros(B, dtype=torch.int32, device="cuda")
    kps = torch.zeros((B, cap, 7), device="cuda"); desc = torch.zeros((B, cap, 32), dtype=torch.uint8, device="cuda")
    stream = torch.cuda.ExternalStream(ex.stream())
    flags = Y.IN_DEVICE | Y.OUT_DEVICE | Y.NO_SYNC
    run = lambda: ex.extract_raw(frames.data_ptr(), B, W, H, W, W * H, flags, cap, counts.data_ptr(), kps.data_ptr(), desc.data_ptr())
    for _ in range(3):
        run()
    ex.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(10):
        run()
    e1.record(stream); ex.sync()
    ms = e0.elapsed_time(e1) / 10
    out[mb] = ms
    print("max_batch %4d: %.3f ms per 512 frames, %.1fk frames/s, kp %d" % (mb, ms, B / ms, int(counts.sum().item())), flush=True)
    ex.close()
print(json.dumps(out))
