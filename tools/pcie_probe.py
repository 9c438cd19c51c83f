"""Pinned host<->device copy bandwidth of the box (the ceiling of bench.py's e2e leg)."""
import torch

dev = torch.device("cuda", 0)
for mb in (64, 256):
    n = mb << 20
    h_in, h_out = torch.empty(n, dtype=torch.uint8).pin_memory(), torch.empty(n, dtype=torch.uint8).pin_memory()
    d_in, d_out = torch.empty(n, dtype=torch.uint8, device=dev), torch.zeros(n, dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    for name in ("h2d", "d2h", "both"):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        s1.wait_event(e0)
        s2.wait_event(e0)
        for _ in range(10):
            if name != "d2h":
                with torch.cuda.stream(s1):
                    d_in.copy_(h_in, non_blocking=True)
            if name != "h2d":
                with torch.cuda.stream(s2):
                    h_out.copy_(d_out, non_blocking=True)
        torch.cuda.current_stream().wait_stream(s1)
        torch.cuda.current_stream().wait_stream(s2)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        print("%-4s %4d MB x10: %7.2f ms  %.1f GB/s per direction" % (name, mb, ms, 10 * n / ms / 1e6))
