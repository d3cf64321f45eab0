"""GPU: parity at BASELINE.json's full sizes, frame by frame against the oracle:
configs[1] all 4096 frames 640x480, configs[3] a 64-frame 1920x1080 blur+noise subset, configs[4] a
10 000-frame subset of the 200-correspondence PnP stress against cv2.solvePnP.  Assertions: Canny and
morphology maps bit-exact and quad lists equal on EVERY frame; every frame within 1e-4 rad / 1 mm of
the oracle's pose.  The per-frame records land in gpurun_out/ (when that directory exists)."""
import json
import os

import numpy as np
import pytest

import parity
from frames import K_for, poses
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _dump(name, summary):
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, name), "w") as f:
            json.dump(summary, f, indent=1)


def _assert_parity(s):
    n = s["frames"]
    assert s["canny_equal"] == n and s["morph_equal"] == n, s
    assert s["quads_equal"] == n, s
    assert s["counts_equal"] == n, s
    assert s["within_tol"] == n, s["offenders"]


def test_config1_all_4096_frames(ctx):
    """BASELINE.json configs[1]: 4096 frames 640x480 rendered on the device (the bench workload, same seed)."""
    import torch

    W, H, N = 640, 480, 4096
    K, D = K_for(W, H), np.zeros(4)
    ps = poses(N, seed=0x4D414E54 + 2)
    buf = torch.empty((N, H, W, 3), dtype=torch.uint8, device="cuda:0")
    ctx.render_grid_device(ps, W, H, K, D, buf.data_ptr())
    torch.cuda.synchronize()
    frames = buf.cpu().numpy()
    del buf
    s, _ = parity.run(ctx, frames, K, D)
    _dump("parity_config1.json", s)
    assert s["status_ok"] == N
    _assert_parity(s)
    assert s["identical"] == N, s["offenders"]


def test_config3_64_frames_1080p_noise_blur(ctx):
    """BASELINE.json configs[3]: 1920x1080 frames with Gaussian blur (sigma 1.5 px) and N(0, 8) noise."""
    import torch

    W, H, N = 1920, 1080, 64
    K, D = K_for(W, H), np.zeros(4)
    clean = torch.empty((N, H, W, 3), dtype=torch.uint8, device="cuda:0")
    noisy = torch.empty_like(clean)
    ctx.render_grid_device(poses(N, seed=0x4D414E54 + 4), W, H, K, D, clean.data_ptr())
    ctx.degrade_device(clean.data_ptr(), noisy.data_ptr(), N, W, H, blur_sigma=1.5, noise_sigma=8.0, seed=4)
    torch.cuda.synchronize()
    frames = noisy.cpu().numpy()
    del clean, noisy
    s, _ = parity.run(ctx, frames, K, D, chunk=32)
    _dump("parity_config3.json", s)
    _assert_parity(s)


def test_config4_pnp_10000_frames(ctx):
    """BASELINE.json configs[4] subset: 10 000 frames x 200 correspondences (SURVEY 8d config 5's generator)
    against cv2.solvePnP(SOLVEPNP_ITERATIVE) and the oracle: 1e-4 rad / 1 mm."""
    from test_pnp import rot_angle
    cv2 = pytest.importorskip("cv2")

    N, M = 10000, 200
    white, red, green = po.make_map()
    mp3 = np.concatenate([white, red, green]).astype(np.float32)
    K = np.array([[320, 0, 320], [0, 320, 240], [0, 0, 1]], np.float32)
    rng = np.random.default_rng(0x4D414E54 + 5)
    ps = poses(N, seed=0x4D414E54 + 6)
    obj = np.empty((N, M, 3), np.float32)
    img = np.empty((N, M, 2), np.float32)
    for i, (Rw, tw) in enumerate(ps):
        ob = mp3[rng.choice(len(mp3), M, replace=False)]
        pc = (Rw.T @ ob.astype(np.float64).T).T - Rw.T @ tw
        obj[i] = ob
        img[i] = np.stack([320 * pc[:, 0] / pc[:, 2] + 320, 320 * pc[:, 1] / pc[:, 2] + 240], 1) + rng.normal(0, 0.5, (M, 2))
    out = ctx.pnp(obj, img, K)
    worst_r = worst_t = 0.0
    bad = []
    for i in range(N):
        ok, r, t = cv2.solvePnP(obj[i].astype(np.float64), img[i].astype(np.float64), K.astype(np.float64), None,
                                flags=cv2.SOLVEPNP_ITERATIVE)
        assert ok and out[i].status == 0
        dr, dt = rot_angle(out[i].rvec[:], r.ravel()), float(np.abs(np.array(out[i].tvec[:]) - t.ravel()).max())
        worst_r, worst_t = max(worst_r, dr), max(worst_t, dt)
        if not (dr < 1e-4 and dt < 1e-3):
            bad.append((i, dr, dt))
    _dump("parity_config4_pnp.json", {"frames": N, "max_drot": worst_r, "max_dt": worst_t, "offenders": bad[:32]})
    assert not bad, bad[:8]
