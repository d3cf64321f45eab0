"""GPU: whole-frame localization through the C ABI against the oracle (same seeded frames):
counts and discrete decisions equal, scores equal, poses within 1e-4 rad / 1 mm."""
import numpy as np
import pytest

from cases import all_cases
from frames import D_UNIT, K_GRID1, K_for, poses
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu


def rot_angle(q1, q2):
    d = abs(float(np.dot(q1, q2)))
    return 2 * np.arccos(min(1.0, d))


def compare(res, tr, otr):
    """Every count, every discrete decision and every integer score equal; pose within 1e-4 rad / 1 mm."""
    assert res.status == otr.status
    assert (tr.n_contours, tr.n_quads_raw, tr.n_quads) == (otr.n_contours, otr.n_quads_raw, otr.n_quads)
    assert (tr.n_hyps_generated, tr.n_hyps_cluster) == (otr.n_hyps_generated, otr.n_hyps_cluster)
    if otr.status != 0:
        return True
    assert tr.first_best_error == otr.first_best.error
    assert tr.optimized_error == otr.optimized.error
    assert list(tr.shift_err) == list(otr.shift_err)
    assert tr.best_rot == otr.best_rot
    assert np.allclose(list(tr.yaw_err), list(otr.yaw_err), rtol=1e-12)
    assert np.abs(np.array(res.pos[:]) - np.array(otr.pos[:])).max() < 1e-3      # 1 mm
    assert rot_angle(np.array(res.quat_xyzw[:]), np.array(otr.quat_xyzw[:])) < 1e-4  # 1e-4 rad
    assert res.error == pytest.approx(otr.error, rel=1e-12)
    assert res.min_yaw_diff == pytest.approx(otr.min_yaw_diff, rel=1e-9, abs=1e-9)
    assert res.n_particles == otr.n_scored
    return True


def test_localize_grid1(ctx):
    img = all_cases()["grid1"]
    res, tr = ctx.localize(img, K_GRID1, D_UNIT, want_trace=True)
    otr, _ = po.localize(img, K_GRID1, D_UNIT)
    assert compare(res[0], tr[0], otr)


def test_localize_synthetic_batch(ctx):
    K, D = K_for(640, 480), np.zeros(4)
    frames = np.stack([po.render_grid(R, t, K, D, 640, 480) for R, t in poses(24, seed=0x4D414E54)])
    res, tr = ctx.localize(frames, K, D, want_trace=True)
    for i in range(len(frames)):
        otr, _ = po.localize(frames[i], K, D)
        compare(res[i], tr[i], otr)


def test_localize_hd(ctx):
    K, D = K_for(1280, 720), np.zeros(4)
    frames = np.stack([po.render_grid(R, t, K, D, 1280, 720) for R, t in poses(3, seed=0x4D414E55)])
    res, tr = ctx.localize(frames, K, D, want_trace=True)
    for i in range(len(frames)):
        otr, _ = po.localize(frames[i], K, D)
        compare(res[i], tr[i], otr)


def test_localize_early_outs(ctx):
    K, D = K_for(64, 48), np.zeros(4)
    blank = np.zeros((2, 48, 64, 3), np.uint8)
    res, _ = ctx.localize(blank, K, D)
    assert res[0].status == 1 and res[1].status == 1  # no quads (src/mantis3.cpp:83-84)


def _blur_noise(img, seed):
    """Config 4's degradation: separable Gaussian blur sigma 1.5 px (7 taps, reflect) + N(0, 8) noise, clipped."""
    rng = np.random.default_rng(seed)
    x = np.arange(-3, 4)
    k = np.exp(-x * x / (2 * 1.5 ** 2))
    k /= k.sum()
    f = img.astype(np.float64)
    for axis in (0, 1):
        pad = [(3, 3) if a == axis else (0, 0) for a in range(3)]
        g = np.pad(f, pad, mode="reflect")
        f = sum(k[i] * np.take(g, np.arange(i, i + img.shape[axis]), axis=axis) for i in range(7))
    return np.clip(np.rint(f + rng.normal(0, 8, f.shape)), 0, 255).astype(np.uint8)


def test_localize_full_hd_with_noise_and_blur(ctx):
    """BASELINE.json configs[3]: 1920x1080 frames with blur + noise; every intermediate count and score equal."""
    K, D = K_for(1920, 1080), np.zeros(4)
    frames = np.stack([_blur_noise(po.render_grid(R, t, K, D, 1920, 1080), 40 + i)
                       for i, (R, t) in enumerate(poses(2, seed=0x4D414E54 + 4))])
    canny, morph = ctx.edges(frames)
    res, tr = ctx.localize(frames, K, D, want_trace=True)
    for i in range(len(frames)):
        oc, om = po.edge_stage(frames[i], 50)
        assert np.array_equal(canny[i], oc) and np.array_equal(morph[i], om)
        otr, _ = po.localize(frames[i], K, D)
        compare(res[i], tr[i], otr)


def _rz(k):
    c, s = np.cos(k * np.pi / 2), np.sin(k * np.pi / 2)
    return np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]])


def _qmat(q):
    x, y, z, w = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def test_localize_full_batch_properties(ctx):
    """BASELINE.json configs[1] at full size (4096 frames 640x480, rendered on the device, device-resident):
    size-independent properties -- determinism, batch invariance, oracle parity on a sample, and the
    render -> localize round trip modulo the grid's symmetry group (quarter turns about z; the reference's
    own accuracy, not ours, sets the tolerance: sigma 0.03 rad perturbations, PoseAdjustment.h:29)."""
    import torch

    import mantis_b200

    W, H, N = 640, 480, 4096
    K, D = K_for(W, H), np.zeros(4)
    ps = poses(N, seed=0x4D414E54 + 2)
    buf = torch.empty((N, H, W, 3), dtype=torch.uint8, device="cuda:0")
    ctx.render_grid_device(ps, W, H, K, D, buf.data_ptr())
    torch.cuda.synchronize()
    ptrs = [buf.data_ptr() + i * H * W * 3 for i in range(N)]
    res, _ = ctx.localize_frames(ctx.make_frames(ptrs, W, H, K, D), mantis_b200.MEM_DEVICE)
    assert all(r.status == 0 for r in res)
    # determinism: the same call again gives the same bytes
    res2, _ = ctx.localize_frames(ctx.make_frames(ptrs, W, H, K, D), mantis_b200.MEM_DEVICE)
    assert bytes(res) == bytes(res2)
    # batch invariance: a slice localized on its own (other chunking, other lanes) gives the same records
    sub, tr_sub = ctx.localize_frames(ctx.make_frames(ptrs[1000:1064], W, H, K, D), mantis_b200.MEM_DEVICE, want_trace=True)
    assert bytes(sub) == bytes(res)[1000 * len(bytes(sub)) // 64:1064 * len(bytes(sub)) // 64]
    # oracle parity on a sample of the slice
    for i in (0, 13, 27, 41, 55, 63):
        frame = buf[1000 + i].cpu().numpy()
        otr, _ = po.localize(frame, K, D)
        compare(sub[i], tr_sub[i], otr)
    # round trip modulo the symmetry group
    ang, dz = [], []
    for i in range(N):
        Rg = _qmat(res[i].quat_xyzw[:])
        ang.append(min(np.arccos(np.clip((np.trace(Rg.T @ (_rz(k) @ ps[i][0])) - 1) / 2, -1, 1)) for k in range(4)))
        dz.append(abs(res[i].pos[2] - ps[i][1][2]))
    assert max(ang) < 0.25 and np.median(ang) < 0.06
    assert max(dz) < 0.3 and np.median(dz) < 0.06


@pytest.mark.parametrize("dist", [False, True])
def test_localize_custom_nonplanar_map(dist):
    """mantis_set_map with a map that leaves the floor plane (z != 0), with and without lens distortion: the
    non-planar / distorted instantiations of the scoring loop against the oracle on the same map."""
    import mantis_b200

    white, red, green = (m.copy() for m in po.make_map())
    rng = np.random.default_rng(7)
    white[::3, 2] = rng.uniform(-0.04, 0.04, len(white[::3]))
    green[:, 2] = 0.02
    K = K_for(640, 480)
    D = D_UNIT if dist else np.zeros(4)
    frames = np.stack([po.render_grid(R, t, K, D, 640, 480) for R, t in poses(3, seed=0x4D414E54 + 9)])
    with mantis_b200.Context(0) as c:
        c.set_map(white, red, green)
        res, tr = c.localize(frames, K, D, want_trace=True)
    for i in range(len(frames)):
        otr, _ = po.localize(frames[i], K, D, maps=(white, red, green))
        compare(res[i], tr[i], otr)
        assert tr[i].first_best_error == otr.first_best.error  # the scores of the cluster never depend on the RPP stop test


def test_localize_odd_size_and_padded_rows(ctx):
    """Ragged input: a frame size that is no multiple of 4 or 32 (byte-wise gray path, partial last words,
    partial tiles), once tightly packed and once with padded rows (stride > 3 W, the 2-D copy path)."""
    import mantis_b200

    W, H = 651, 487
    K, D = K_for(W, H), np.zeros(4)
    frames = np.stack([po.render_grid(R, t, K, D, W, H) for R, t in poses(2, seed=0x4D414E54 + 11)])
    res, tr = ctx.localize(frames, K, D, want_trace=True)
    stride = 3 * W + 5
    padded = np.zeros((len(frames), H, stride), np.uint8)
    padded[:, :, :3 * W] = frames.reshape(len(frames), H, 3 * W)
    fr = ctx.make_frames([padded[i] for i in range(len(frames))], W, H, K, D, stride=stride)
    res_p, _ = ctx.localize_frames(fr, mantis_b200.MEM_HOST)
    assert bytes(res_p) == bytes(res)
    for i in range(len(frames)):
        otr, _ = po.localize(frames[i], K, D)
        compare(res[i], tr[i], otr)
        assert (tr[i].n_contours, tr[i].n_quads) == (otr.n_contours, otr.n_quads)


def test_localize_capacity_limits_are_reported(ctx):
    """Frames beyond the per-frame capacities (1024 contours, 256 quads; the reference has none) must come back
    as MANTIS_FRAME_OVERFLOW with the flag that says why, never as a pose; a frame of pure noise (143 contours,
    no 4-gon survives) must match the oracle."""
    K, D = K_for(640, 480), np.zeros(4)
    rng = np.random.default_rng(3)
    noise = rng.integers(0, 256, (480, 640, 3), dtype=np.uint8)
    fine, coarse = np.zeros((480, 640, 3), np.uint8), np.zeros((480, 640, 3), np.uint8)
    for y in range(0, 480, 12):
        for x in range(0, 640, 12):
            fine[y + 2:y + 10, x + 2:x + 10] = 255      # 2 x 1060 borders + the frame's
    for y in range(0, 480, 24):
        for x in range(0, 640, 24):
            coarse[y + 3:y + 21, x + 3:x + 21] = 255    # 520 squares
    res, tr = ctx.localize(np.stack([noise, fine, coarse]), K, D, want_trace=True)
    otr, _ = po.localize(noise, K, D)
    assert compare(res[0], tr[0], otr)
    for i in (1, 2):
        assert res[i].status == 4 and tr[i].flags != 0
        assert res[i].error == 0 and all(v == 0 for v in res[i].pos[:])


def test_localize_small_grid_keeps_reference_order():
    """grid_size = 4 gives 16 cell shifts: getBestNHypotheses(20) then returns the list UNSORTED
    (HypothesisEvaluation.h:492), so hyps.back() is the last shift, not the best."""
    import mantis_b200

    K, D = K_for(640, 480), np.zeros(4)
    frames = np.stack([po.render_grid(R, t, K, D, 640, 480) for R, t in poses(4, seed=0x4D414E54 + 13)])
    cfg = mantis_b200.binding.default_config()
    cfg.grid_size = 4
    prm = po.default_params()
    prm.grid_size = 4
    with mantis_b200.Context(0, cfg) as c:
        res, tr = c.localize(frames, K, D, want_trace=True)
    for i in range(len(frames)):
        otr, _ = po.localize(frames[i], K, D, params=prm)
        assert list(tr[i].shift_err[:16]) == list(otr.shift_err[:16])
        assert tr[i].best_rot == otr.best_rot
        assert np.allclose(list(tr[i].yaw_err), list(otr.yaw_err), rtol=1e-12)
        assert np.abs(np.array(res[i].pos[:]) - np.array(otr.pos[:])).max() < 1e-3
        assert rot_angle(np.array(res[i].quat_xyzw[:]), np.array(otr.quat_xyzw[:])) < 1e-4
        assert res[i].error == pytest.approx(otr.error, rel=1e-12)


def test_localize_rejects_skewed_camera_matrix(ctx):
    import mantis_b200

    K = K_for(64, 48).copy()
    K[0, 1] = 0.5
    with pytest.raises(mantis_b200.MantisError, match="skew"):
        ctx.localize(np.zeros((1, 48, 64, 3), np.uint8), K, np.zeros(4))


def test_localize_async_equals_blocking_call(ctx):
    """mantis_localize_batch_async + mantis_wait: the same records as the blocking call, for device-resident frames
    (several chunks queued back to back, results into pinned memory) and for host frames."""
    import ctypes

    import torch

    import mantis_b200

    W, H, N = 640, 480, 300
    K, D = K_for(W, H), np.zeros(4)
    buf = torch.empty((N, H, W, 3), dtype=torch.uint8, device="cuda:0")
    ctx.render_grid_device(poses(N, seed=0x4D414E54 + 14), W, H, K, D, buf.data_ptr())
    torch.cuda.synchronize()
    dframes = ctx.make_frames([buf.data_ptr() + i * H * W * 3 for i in range(N)], W, H, K, D)
    ref, _ = ctx.localize_frames(dframes, mantis_b200.MEM_DEVICE)
    pinned = torch.zeros(N * ctypes.sizeof(mantis_b200.Result), dtype=torch.uint8, pin_memory=True)
    out = (mantis_b200.Result * N).from_address(pinned.data_ptr())
    ctx.localize_frames_async(dframes, mantis_b200.MEM_DEVICE, out)
    ctx.wait()
    assert bytes(out) == bytes(ref)
    host = buf[:40].cpu().numpy()
    hframes = ctx.make_frames([host[i] for i in range(40)], W, H, K, D)
    out2 = (mantis_b200.Result * 40)()
    ctx.localize_frames_async(hframes, mantis_b200.MEM_HOST, out2)
    ctx.wait()
    assert bytes(out2) == bytes(ref)[:40 * ctypes.sizeof(mantis_b200.Result)]


_SHIFT_CHILD = r"""
import ctypes, sys
import numpy as np, torch
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[1] + "/tests")
import mantis_b200
from frames import K_for, poses
W, H, N = 640, 480, 200
K, D = K_for(W, H), np.zeros(4)
with mantis_b200.Context(0) as ctx:
    buf = torch.empty((N, H, W, 3), dtype=torch.uint8, device="cuda:0")
    ctx.render_grid_device(poses(N, seed=0x4D414E54 + 21), W, H, K, D, buf.data_ptr())
    torch.cuda.synchronize()
    frames = ctx.make_frames([buf.data_ptr() + i * H * W * 3 for i in range(N)], W, H, K, D)
    res, tr = ctx.localize_frames(frames, mantis_b200.MEM_DEVICE, want_trace=True)
    open(sys.argv[2], "wb").write(bytes(res) + bytes(tr))
"""


def test_shift_hypotheses_through_the_lattice_equal_direct_scoring(tmp_path):
    """k_search scores the 81 cell shifts from a table of the distinct points p - shift (search.cu: score_points);
    MANTIS_DIRECT_SHIFTS=1 projects every shift on its own, as the reference does.  Results and traces (all 81 shift
    errors, the kept 20, yaw errors, pose) of 200 frames must be the same bytes.  (The variable is read once per
    process, hence the two child processes.)"""
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for name, extra in (("table", {}), ("direct", {"MANTIS_DIRECT_SHIFTS": "1"})):
        path = str(tmp_path / (name + ".bin"))
        env = {k: v for k, v in os.environ.items() if k != "MANTIS_DIRECT_SHIFTS"}
        env.update(extra)
        subprocess.run([sys.executable, "-c", _SHIFT_CHILD, root, path], check=True, env=env, timeout=600)
        outs.append(open(path, "rb").read())
    assert len(outs[0]) > 200 * 64 and outs[0] == outs[1]
