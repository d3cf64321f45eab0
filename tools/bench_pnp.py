"""Config 5: pose-solver-only stress -- N frames x 200 correspondences through mantis_pnp_batch with
device-resident inputs, timed with CUDA events on the library's stream.
Usage: python tools/bench_pnp.py [N M iters]"""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import mantis_b200
from mantis_b200.binding import PnpResult


def synth(N, M, dev, seed=0x4D414E54 + 5):
    """200 of the map points per frame, config-2 pose distribution, pinhole 320/320/240, 0.5 px noise."""
    g = torch.Generator(device=dev).manual_seed(seed)
    line = torch.arange(37, device=dev, dtype=torch.float64) * 0.08 - 1.44
    cross = torch.arange(10, device=dev, dtype=torch.float64) * 0.32 - 1.44
    a = torch.stack(torch.meshgrid(cross, line, indexing="ij"), -1).reshape(-1, 2)
    mp = torch.unique(torch.cat([a, a.flip(1)]), dim=0)                       # 640 unique points
    idx = torch.rand((N, len(mp)), device=dev, generator=g).argsort(1)[:, :M]
    obj = torch.cat([mp[idx], torch.zeros((N, M, 1), device=dev, dtype=torch.float64)], 2)
    yaw = torch.rand(N, device=dev, generator=g, dtype=torch.float64) * 2 * np.pi
    roll, pitch = (torch.randn(N, device=dev, generator=g, dtype=torch.float64) * 0.1 for _ in range(2))
    def rot(axis, ang):
        c, s, z, o = torch.cos(ang), torch.sin(ang), torch.zeros_like(ang), torch.ones_like(ang)
        m = {"x": [o, z, z, z, c, -s, z, s, c], "y": [c, z, s, z, o, z, -s, z, c], "z": [c, -s, z, s, c, z, z, z, o]}[axis]
        return torch.stack(m, -1).reshape(-1, 3, 3)
    flip = torch.diag(torch.tensor([1.0, -1.0, -1.0], device=dev, dtype=torch.float64))
    Rw = rot("z", yaw) @ flip @ rot("y", pitch) @ rot("x", roll)                # camera -> world
    tw = torch.stack([torch.rand(N, device=dev, generator=g, dtype=torch.float64) * 2 - 1,
                      torch.rand(N, device=dev, generator=g, dtype=torch.float64) * 2 - 1,
                      torch.rand(N, device=dev, generator=g, dtype=torch.float64) * 1.2 + 0.8], 1)
    pc = torch.einsum("nji,nmj->nmi", Rw, obj - tw[:, None, :])                # world -> camera
    uv = torch.stack([320 * pc[..., 0] / pc[..., 2] + 320, 320 * pc[..., 1] / pc[..., 2] + 240], -1)
    uv = uv + torch.randn(uv.shape, device=dev, generator=g, dtype=torch.float64) * 0.5
    return obj.float().contiguous(), uv.float().contiguous(), Rw, tw


def main():
    N, M, iters = (int(x) for x in (sys.argv[1:4] + ["1000000", "200", "3"][len(sys.argv) - 1:]))
    dev = torch.device("cuda:0")
    chunk = 1 << 17
    parts = [synth(min(chunk, N - s), M, dev, seed=1234 + s) for s in range(0, N, chunk)]
    obj = torch.cat([p[0] for p in parts]); img = torch.cat([p[1] for p in parts])
    Rw = torch.cat([p[2] for p in parts]); tw = torch.cat([p[3] for p in parts])
    del parts
    K = torch.tensor([320, 0, 320, 0, 320, 240, 0, 0, 1], dtype=torch.float32, device=dev)
    out = torch.empty((N, ctypes.sizeof(PnpResult)), dtype=torch.uint8, device=dev)
    ctx = mantis_b200.Context(0)
    stream = torch.cuda.ExternalStream(ctx.stream)
    torch.cuda.synchronize()
    run = lambda: ctx.pnp_device(obj.data_ptr(), img.data_ptr(), N, M, K.data_ptr(), out.data_ptr())
    for _ in range(2):
        run()
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); run(); e1.record(stream)
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = min(ts)
    res = np.frombuffer(out.cpu().numpy().tobytes(), dtype=np.dtype([("rvec", "3f8"), ("tvec", "3f8"), ("err", "f8"), ("passes", "i4"), ("status", "i4")]))
    # ground-truth check: camera position error (the estimate is world->camera: cam origin = -R^T t)
    import cv2
    sub = np.linspace(0, N - 1, 2000).astype(int)
    pos_err = []
    for i in sub:
        R, _ = cv2.Rodrigues(res["rvec"][i])
        pos_err.append(np.abs(-R.T @ res["tvec"][i] - tw[i].cpu().numpy()).max())
    bytes_per_frame = M * (12 + 8) + 64
    print(f"pnp {N} frames x {M} correspondences: {ms:.2f} ms best  {N / ms * 1e3 / 1e6:.2f} M frames/s  "
          f"{N * bytes_per_frame / ms / 1e6:.1f} GB/s algorithmic  mean passes {res['passes'].mean():.2f}  ok {np.mean(res['status'] == 0):.4f}  "
          f"reproj RMS {np.sqrt((res['err'] ** 2).mean() / (2 * M)):.3f} px  median |cam pos err| {np.median(pos_err) * 1e3:.2f} mm")


if __name__ == "__main__":
    main()
