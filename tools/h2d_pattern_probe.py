import torch
dev = torch.device("cuda", 0)
fr = [torch.empty(256 * 752 * 480, dtype=torch.uint8).pin_memory() for _ in range(2)]
dfr = [torch.empty_like(f, device=dev) for f in fr]
sizes = [2048, 6144000, 6144000, 6144000, 6144000, 1536000, 1536000, 49152000]
mp = [torch.empty(s, dtype=torch.uint8).pin_memory() for s in sizes]
dmp = [torch.empty_like(m, device=dev) for m in mp]
s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
def step(two_streams):
    with torch.cuda.stream(s2 if two_streams else s1):
        for h, d in zip(mp, dmp):
            d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s1):
        for h, d in zip(fr, dfr):
            d.copy_(h, non_blocking=True)
for two in (True, False):
    for _ in range(3):
        step(two)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); s1.wait_event(e0); s2.wait_event(e0)
    for _ in range(10):
        step(two)
    torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)
    e1.record(); torch.cuda.synchronize()
    tot = sum(sizes) + 2 * 256 * 752 * 480
    ms = e0.elapsed_time(e1) / 10
    print("two streams" if two else "one stream", "%.3f ms per step, %.1f GB/s" % (ms, tot / ms / 1e6))
