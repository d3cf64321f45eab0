"""One pass of every kernel worth an ncu section, at modest sizes: the localization chain (1024 frames 640x480,
device-resident), the line-voting stage (64 frames) and the PnP solver (65 536 frames x 200 correspondences).
  ncu --set full --clock-control none --import-source on -o gpurun_out/r2_full python tools/profile_all.py"""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tools")):
    sys.path.insert(0, p)
import mantis_b200  # noqa: E402
from bench_pnp import synth  # noqa: E402
from frames import K_for, poses  # noqa: E402
from mantis_b200.binding import PnpResult  # noqa: E402


def main():
    W, H, N = 640, 480, int(os.environ.get("PROFILE_FRAMES", "1024"))
    K, D = K_for(W, H), np.zeros(4)
    dev = torch.device("cuda:0")
    ctx = mantis_b200.Context(0)
    buf = torch.empty((N, H, W, 3), dtype=torch.uint8, device=dev)
    ctx.render_grid_device(poses(N, seed=0x4D414E54 + 2), W, H, K, D, buf.data_ptr())
    frames = ctx.make_frames([buf.data_ptr() + i * H * W * 3 for i in range(N)], W, H, K, D)
    res, _ = ctx.localize_frames(frames, mantis_b200.MEM_DEVICE)
    print("localize ok", sum(r.status == 0 for r in res), "/", N)
    lines = ctx.hough(buf[:64].cpu().numpy(), threshold=120)
    print("hough peaks", [l.n_peaks for l in lines][:4])
    M, NP = 200, 65536
    obj, img, _, _ = synth(NP, M, dev)
    Kt = torch.tensor([320, 0, 320, 0, 320, 240, 0, 0, 1], dtype=torch.float32, device=dev)
    out = torch.empty((NP, ctypes.sizeof(PnpResult)), dtype=torch.uint8, device=dev)
    ctx.pnp_device(obj.data_ptr(), img.data_ptr(), NP, M, Kt.data_ptr(), out.data_ptr())
    torch.cuda.synchronize()
    print("pnp done")
    ctx.close()


if __name__ == "__main__":
    main()
