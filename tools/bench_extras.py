"""Extra keys of bench.py's JSON line (N = 1 only): the other BASELINE.json configs and the parity sample, each
measured in the same run with CUDA events on the library's stream (device-resident inputs, 1 warm-up + `reps` timed
calls, best time).  Frames are rendered on the device by the library's own frame source.

  hd720_frames_per_s          configs[2]'s resolution: 1280x720, one camera
  rig_4x720p_rigs_per_s       configs[2]: rigs of 4 cameras 1280x720, joint localization (mantis_localize_rig_batch)
  hd1080_noise_frames_per_s   configs[3]: 1920x1080 with Gaussian blur (sigma 1.5 px) + N(0, 8) noise
  d_nonzero_frames_per_s      configs[1] with the reference camera's fisheye coefficients (test/unit_test.cpp:144)
  pnp_frames_per_s            configs[4]: 1 M frames x 200 correspondences through mantis_pnp_batch
  parity_sample               frames of the bench workload localized by the CUDA path AND the oracle (outside any timer)
"""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)


def _best_ms(stream, fn, reps):
    import torch

    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        fn()
        e1.record(stream)
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts)


def localize_rate(ctx, W, H, N, D, seed, reps=2, degrade=False):
    """frames/s of mantis_localize_batch on N device-resident rendered frames W x H."""
    import torch

    import mantis_b200
    from frames import K_for, poses

    K = K_for(W, H)
    dev = torch.device("cuda", ctx.device)
    buf = torch.empty((N, H, W, 3), dtype=torch.uint8, device=dev)
    ctx.render_grid_device(poses(N, seed=seed), W, H, K, D, buf.data_ptr())
    if degrade:
        clean = buf
        buf = torch.empty_like(clean)
        ctx.degrade_device(clean.data_ptr(), buf.data_ptr(), N, W, H, blur_sigma=1.5, noise_sigma=8.0, seed=seed & 0xffff)
        del clean
    torch.cuda.synchronize()
    frames = ctx.make_frames([buf.data_ptr() + i * H * W * 3 for i in range(N)], W, H, K, D)
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
    res = [None]

    def run():
        res[0], _ = ctx.localize_frames(frames, mantis_b200.MEM_DEVICE)
    ms = _best_ms(stream, run, reps)
    ok = sum(1 for r in res[0] if r.status == 0)
    return {"value": N / (ms * 1e-3), "frames": N, "ms": ms, "frames_ok": ok}


def rig_rate(ctx, W, H, n_rigs, seed, reps=2):
    """rigs/s of mantis_localize_rig_batch: 4 cameras (down, 30 degrees fore / aft / side) per rig, device-resident."""
    import torch

    import mantis_b200
    from frames import K_for, poses, rig_cam_to_base, rig_extrinsics

    K, D = K_for(W, H), np.zeros(4)
    ext, b2c = rig_extrinsics(), rig_cam_to_base()
    C = len(ext)
    cam_poses = [(Rb @ Rc, tb + Rb @ tc) for Rb, tb in poses(n_rigs, seed=seed) for Rc, tc in ext]
    dev = torch.device("cuda", ctx.device)
    buf = torch.empty((n_rigs * C, H, W, 3), dtype=torch.uint8, device=dev)
    ctx.render_grid_device(cam_poses, W, H, K, D, buf.data_ptr())
    torch.cuda.synchronize()
    frames = ctx.make_frames([buf.data_ptr() + i * H * W * 3 for i in range(n_rigs * C)], W, H, K, D)
    for i in range(n_rigs * C):
        frames[i].cam_to_base[:] = b2c[i % C].tolist()
        frames[i].camera_id = i % C
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
    res = [None]

    def run():
        res[0], _ = ctx.localize_rig_frames(frames, n_rigs, C, mantis_b200.MEM_DEVICE)
    ms = _best_ms(stream, run, reps)
    ok = sum(1 for r in res[0] if r.status == 0)
    return {"value": n_rigs / (ms * 1e-3), "rigs": n_rigs, "cameras_per_rig": C, "ms": ms, "rigs_ok": ok,
            "camera_frames_per_s": n_rigs * C / (ms * 1e-3)}


def pnp_rate(ctx, N=1_000_000, M=200, reps=2):
    import torch

    from bench_pnp import synth
    from mantis_b200.binding import PnpResult

    dev = torch.device("cuda", ctx.device)
    chunk = 1 << 17
    parts = [synth(min(chunk, N - s), M, dev, seed=1234 + s)[:2] for s in range(0, N, chunk)]
    obj = torch.cat([p[0] for p in parts])
    img = torch.cat([p[1] for p in parts])
    del parts
    Kt = torch.tensor([320, 0, 320, 0, 320, 240, 0, 0, 1], dtype=torch.float32, device=dev)
    out = torch.empty((N, ctypes.sizeof(PnpResult)), dtype=torch.uint8, device=dev)
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
    torch.cuda.synchronize()
    ms = _best_ms(stream, lambda: ctx.pnp_device(obj.data_ptr(), img.data_ptr(), N, M, Kt.data_ptr(), out.data_ptr()), reps)
    res = np.frombuffer(out.cpu().numpy().tobytes(),
                        dtype=np.dtype([("rvec", "3f8"), ("tvec", "3f8"), ("err", "f8"), ("passes", "i4"), ("status", "i4")]))
    bytes_per_frame = M * (12 + 8) + 64
    return {"value": N / (ms * 1e-3), "frames": N, "correspondences": M, "ms": ms,
            "algorithmic_gb_per_s": N * bytes_per_frame / (ms * 1e-3) / 1e9, "mean_passes": float(res["passes"].mean()),
            "ok_fraction": float(np.mean(res["status"] == 0)),
            "reproj_rms_px": float(np.sqrt((res["err"] ** 2).mean() / (2 * M)))}


def parity_sample(ctx, device_frames, W, H, K, D, n):
    """The first n frames of the bench workload through the CUDA path and through the oracle (one worker per host
    core): how many are identical in every decision and score, how many within 1e-4 rad / 1 mm."""
    import parity

    host = device_frames[:n].cpu().numpy()
    s, _ = parity.run(ctx, host, K, D, want_edges=True, want_quads=True, chunk=256)
    return {"frames": s["frames"], "identical": s["identical"], "within_tol": s["within_tol"],
            "edges_bit_exact": min(s.get("canny_equal", 0), s.get("morph_equal", 0)), "quads_equal": s.get("quads_equal"),
            "max_dpos_m": s["max_dpos"], "max_drot_rad": s["max_drot"], "tolerance": "1e-4 rad, 1e-3 m (BASELINE.json north_star)"}
