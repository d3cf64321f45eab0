"""Times the device-resident image stage (mantis_edges_batch, MEM_DEVICE) with CUDA events on the
library's stream.  Usage: python tools/bench_edges.py [W H N iters]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import mantis_b200
from cases import blocky


def main():
    W, H, N, iters = (int(a) for a in (sys.argv[1:5] + ["640", "480", "4096", "5"][len(sys.argv) - 1:]))
    rng = np.random.default_rng(0)
    base = np.stack([blocky(rng, H, W, 24 + 8 * i) for i in range(8)])
    dev = torch.device("cuda:0")
    frames = torch.from_numpy(base).to(dev).repeat((N + 7) // 8, 1, 1, 1)[:N].contiguous()
    morph = torch.empty((N, H, W), dtype=torch.uint8, device=dev)
    ctx = mantis_b200.Context(0)
    stream = torch.cuda.ExternalStream(ctx.stream)
    torch.cuda.synchronize()
    for _ in range(3):
        ctx.edges_device(frames.data_ptr(), N, W, H, d_morph=morph.data_ptr())
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        ctx.edges_device(frames.data_ptr(), N, W, H, d_morph=morph.data_ptr())
        e1.record(stream)
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = min(ts)
    gb = N * W * H * 4 / 1e9
    print(f"{W}x{H} x{N}: {ms:.3f} ms best ({np.mean(ts):.3f} mean)  {N / ms * 1e3:.0f} frames/s  "
          f"{gb / ms * 1e3:.1f} GB/s algorithmic (4 B/px)")


if __name__ == "__main__":
    main()
