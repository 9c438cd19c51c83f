"""Small fixed workload for ncu: N steps of the device-resident extraction step of bench.py."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402
import ygzorb_ctypes as Y  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    frames = bench.make_frames(B, 0)
    dev = torch.device("cuda", 0)
    dev_in = torch.from_numpy(frames).to(dev)
    ex = Y.Extractor(bench.NFEAT, bench.SCALE, bench.LEVELS, bench.INI_TH, bench.MIN_TH, max_batch=B)
    cap = ex.max_keypoints(bench.W, bench.H)
    d_counts = torch.zeros(B, dtype=torch.int32, device=dev)
    d_kps = torch.zeros((B, cap, 7), dtype=torch.float32, device=dev)
    d_desc = torch.zeros((B, cap, 32), dtype=torch.uint8, device=dev)
    for _ in range(steps):
        ex.extract_raw(dev_in.data_ptr(), B, bench.W, bench.H, bench.W, bench.W * bench.H,
                       Y.IN_DEVICE | Y.OUT_DEVICE, cap, d_counts.data_ptr(), d_kps.data_ptr(), d_desc.data_ptr())
    print("keypoints", int(d_counts.sum().item()))


if __name__ == "__main__":
    main()
