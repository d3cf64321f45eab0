"""GPU: edge mask and quad extraction (through the C ABI) against the oracle: mask bit-exact,
contour / 4-gon counts equal, quad vertex lists equal in the reference's order, stretched test
points bit-exact."""
import numpy as np
import pytest

from cases import all_cases
from frames import D_UNIT, K_GRID1, K_for, poses
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu

CASES = ["grid1", "lena", "blocky_720x1280", "blocky_1080x1920", "odd_33x129", "ramps_480x640", "tiny_3x5",
         "row_1x64", "col_70x1", "noise_123x77"]


@pytest.mark.parametrize("name", CASES)
def test_mask_matches_oracle(ctx, name):
    img = all_cases()[name]
    mask = ctx.edge_mask(img)[0]
    canny, _ = po.edge_stage(img, 50)
    assert np.array_equal(mask, po.clean_mask(canny)), np.argwhere(mask != po.clean_mask(canny))[:5]


def check_quads(fq, tr):
    assert fq.flags == 0
    assert (fq.n_contours, fq.n_quads_raw, fq.n_quads) == (tr.n_contours, tr.n_quads_raw, tr.n_quads)
    for i in range(tr.n_quads):
        assert list(fq.corners[i]) == list(tr.quad_corners[i]), i
        assert list(fq.test_px[i]) == list(tr.quad_test_px[i]), i


@pytest.mark.parametrize("name", CASES)
def test_quads_match_oracle(ctx, name):
    img = all_cases()[name]
    fq = ctx.quads(img)[0]
    tr, _ = po.localize(img, K_GRID1, D_UNIT)
    check_quads(fq, tr)


def test_quads_synthetic_batch(ctx):
    K, D = K_for(640, 480), np.zeros(4)
    frames = np.stack([po.render_grid(R, t, K, D, 640, 480) for R, t in poses(12, seed=77)])
    out = ctx.quads(frames)
    total = 0
    for i in range(len(frames)):
        tr, _ = po.localize(frames[i], K, D)
        check_quads(out[i], tr)
        total += tr.n_quads
    assert total > 100


def test_quads_nested_and_thin_shapes(ctx):
    """Islands in holes, 1-px lines shared by two borders, shapes touching the frame."""
    img = np.zeros((120, 160, 3), np.uint8)
    img[10:100, 10:150] = 255
    img[25:85, 30:130] = 0
    img[40:70, 50:110] = 200
    img[50:60, 70:90] = 0
    img[0:8, 0:40] = 255
    img[110:120, 100:160] = 180
    fq = ctx.quads(img)[0]
    tr, _ = po.localize(img, K_for(160, 120), np.zeros(4))
    check_quads(fq, tr)
    assert tr.n_contours >= 6


def test_device_renderer_matches_oracle(ctx):
    """The CUDA ray-caster against the oracle's: identical up to boundary pixels where the two
    libm's tan/atan differ in the last bit."""
    import torch

    K, D = K_for(640, 480), np.zeros(4)
    ps = poses(4, seed=99)
    buf = torch.empty((len(ps), 480, 640, 3), dtype=torch.uint8, device="cuda:0")
    ctx.render_grid_device(ps, 640, 480, K, D, buf.data_ptr())
    got = buf.cpu().numpy()
    for i, (R, t) in enumerate(ps):
        ref = po.render_grid(R, t, K, D, 640, 480)
        diff = (got[i] != ref).any(axis=2).mean()
        assert diff < 1e-4, diff
        assert (ref > 200).any()


def _blob_images(n, h, w, seed):
    """Smoothed-noise blobs at several scales: nested holes, spirals of thin walls, frame contacts."""
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        f = rng.random((h, w))
        k = 2 + i % 7
        for _ in range(2):  # box blur by cumulative sums
            c = np.cumsum(np.pad(f, ((k, k), (0, 0)), mode="wrap"), axis=0)
            f = c[2 * k:] - c[:-2 * k]
            c = np.cumsum(np.pad(f, ((0, 0), (k, k)), mode="wrap"), axis=1)
            f = c[:, 2 * k:] - c[:, :-2 * k]
        f = (f - f.min()) / (np.ptp(f) + 1e-12)
        g = (np.floor(f * (3 + i % 3)) % 2 * 255).astype(np.uint8)
        out.append(np.repeat(g[:, :, None], 3, axis=2))
    return np.stack(out)


def test_quads_random_blobs_match_oracle(ctx):
    frames = _blob_images(40, 150, 210, seed=5)
    K, D = K_for(210, 150), np.zeros(4)
    out = ctx.quads(frames)
    contours = 0
    for i in range(len(frames)):
        tr, _ = po.localize(frames[i], K, D)
        check_quads(out[i], tr)
        contours += tr.n_contours
    assert contours > 400


def test_quad_routes_agree(ctx, monkeypatch):
    """The piece-parallel tracer and the sequential-walk kernels it falls back to give the same frames."""
    K, D = K_for(640, 480), np.zeros(4)
    frames = np.stack([po.render_grid(R, t, K, D, 640, 480) for R, t in poses(6, seed=91)])
    rng = np.random.default_rng(2)
    frames[3] = np.clip(frames[3].astype(np.int16) + rng.integers(-60, 60, frames[3].shape), 0, 255).astype(np.uint8)
    a = ctx.quads(frames)
    monkeypatch.setenv("MANTIS_QUADS_SEQUENTIAL", "1")
    b = ctx.quads(frames)
    for fa, fb in zip(a, b):
        assert (fa.flags, fa.n_contours, fa.n_quads_raw, fa.n_quads) == (fb.flags, fb.n_contours, fb.n_quads_raw, fb.n_quads)
        for i in range(fa.n_quads):
            assert list(fa.corners[i]) == list(fb.corners[i])
            assert list(fa.test_px[i]) == list(fb.test_px[i])
