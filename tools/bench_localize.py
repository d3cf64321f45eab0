"""Quick device-resident timing of mantis_localize_batch on GPU-rendered frames.
Usage: python tools/bench_localize.py [W H N iters]"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import mantis_b200
from frames import K_for, poses


def main():
    W, H, N, iters = (int(a) for a in (sys.argv[1:5] + ["640", "480", "1024", "3"][len(sys.argv) - 1:]))
    K, D = K_for(W, H), np.zeros(4)
    ctx = mantis_b200.Context(0)
    buf = torch.empty((N, H, W, 3), dtype=torch.uint8, device="cuda:0")
    ctx.render_grid_device(poses(N, seed=1234), W, H, K, D, buf.data_ptr())
    frames = ctx.make_frames([buf.data_ptr() + i * H * W * 3 for i in range(N)], W, H, K, D)
    torch.cuda.synchronize()
    res, _ = ctx.localize_frames(frames, mantis_b200.MEM_DEVICE)
    st = np.bincount([r.status for r in res], minlength=5)
    print("status histogram", st.tolist(), "mean quads", np.mean([r.n_quads for r in res]), "mean hyps",
          np.mean([r.n_hyps for r in res]), "mean scored", np.mean([r.n_particles for r in res]))
    ctx.set_profiling(True)
    import ctypes
    clk = (ctypes.c_longlong * 16)()
    ctx.localize_frames(frames, mantis_b200.MEM_DEVICE)  # untimed profiled pass (single-lane scratch)
    ctx.stage_ms()
    ctx._lib.mantis_debug_clocks(ctx._h, None)
    ts = []
    for _ in range(iters):
        t0 = time.perf_counter()
        ctx.localize_frames(frames, mantis_b200.MEM_DEVICE)
        ts.append(time.perf_counter() - t0)
    ms = ctx.stage_ms()
    ctx._lib.mantis_debug_clocks(ctx._h, clk)
    tot = sum(clk) or 1
    per = [round(v / (N * iters)) for v in clk]
    print('search kernel cycles/frame (first score, refinement, shifts, yaw windows):', per[:4])
    print('edge-link kernel cycles/frame (load, hysteresis, canny store, 3 morph passes, morph store, mask chain):', [per[4], per[5], per[8], per[9], per[6], per[7]])
    print('ObjPose steps/frame spent in runs of <16, <64, <256, <1024, <4096, more steps:', per[10:16])
    print({k: round(v / iters, 2) for k, v in ms.items()})
    print(f"{W}x{H} x{N}: best {min(ts)*1e3:.1f} ms  -> {N/min(ts):.0f} frames/s")


if __name__ == "__main__":
    main()
