"""Line voting (row a6'): the oracle against cv2.HoughLines' golden lists (CPU), and the CUDA stage
against the oracle through the C ABI (GPU).  The reference has no Hough transform: the pin is the
OpenCV it links, not a reference call site."""
import hashlib
import json
import os

import numpy as np
import pytest

from cases import all_cases
from frames import K_for, poses
from oracle import pyoracle as po

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "hough_golden.json")))


def digest(lines):
    a = np.ascontiguousarray(np.asarray(lines, np.float32).reshape(-1, 2))
    return hashlib.md5(a.tobytes()).hexdigest()


@pytest.mark.parametrize("name", sorted(GOLD["cases"]))
def test_oracle_lines_match_cv2_golden(name):
    g = GOLD["cases"][name]
    canny, _ = po.edge_stage(all_cases()[name], 50)
    lines = po.hough_lines(canny, g["threshold"])
    assert len(lines) == g["n_lines"]
    assert digest(lines[:, :2]) == g["lines_md5"]


def test_oracle_lines_match_cv2_live():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(3)
    for shape, thr in [((61, 93), 12), ((200, 333), 30)]:
        e = (rng.random(shape) < 0.08).astype(np.uint8) * 255
        e[shape[0] // 2, :] = 255
        e[:, shape[1] // 3] = 255
        ref = cv2.HoughLines(e, 1, np.pi / 180, thr)
        ref = np.zeros((0, 2), np.float32) if ref is None else ref.reshape(-1, 2)
        got = po.hough_lines(e, thr)
        assert len(got) == len(ref) and np.array_equal(got[:, :2].astype(np.float32), ref)


def test_oracle_clusters_and_crossings_on_a_clean_grid():
    """Two pencils of 5 + 4 lines drawn 3 px wide: every line becomes one cluster, 20 crossings."""
    e = np.zeros((240, 320), np.uint8)
    for x in (40, 100, 160, 220, 280):
        e[:, x - 1:x + 2] = 255
    for y in (30, 90, 150, 210):
        e[y - 1:y + 2, :] = 255
    lines = po.hough_lines(e, 150)
    cl = po.hough_clusters(lines, 320, 240)
    pts, pen = po.hough_intersections(cl, 320, 240)
    assert len(cl) == 9 and sorted(pen.tolist()) in ([0] * 5 + [1] * 4, [0] * 4 + [1] * 5)
    assert len(pts) == 20
    want = np.array([[x, y] for x in (40, 100, 160, 220, 280) for y in (30, 90, 150, 210)], np.float64)
    d = np.abs(pts[:, None, :].astype(np.float64) - want[None]).max(axis=2)  # plateau peaks sit on the line's first row
    assert (d.min(axis=1) <= 1.5).all() and sorted(d.argmin(axis=1).tolist()) == list(range(20))


def oracle_frame(img, thr, use_morph=False):
    canny, morph = po.edge_stage(img, 50)
    h, w = canny.shape
    lines = po.hough_lines(morph if use_morph else canny, thr, max_lines=1 << 16)
    n_peaks = len(lines)
    lines = lines[:1024]
    cl = po.hough_clusters(lines, w, h)
    pts, _ = po.hough_intersections(cl, w, h)
    return n_peaks, lines, cl, pts


def check_frame(fl, ref):
    n_peaks, lines, cl, pts = ref
    assert fl.n_peaks == n_peaks and fl.flags == (1 if n_peaks > 4096 else 0)
    assert fl.n_lines == len(lines)
    if n_peaks > 4096:  # peak buffer overflow: the kept subset is unspecified (flag says so)
        return
    got = np.array([[fl.lines[i].rho, fl.lines[i].theta, fl.lines[i].votes] for i in range(fl.n_lines)]).reshape(-1, 3)
    assert np.array_equal(got, lines), np.argwhere(got != lines)[:5]
    assert fl.n_clusters == len(cl)
    gc = np.array([[c.rho, c.theta, c.votes, c.members] for c in fl.clusters[:fl.n_clusters]]).reshape(-1, 4)
    assert np.array_equal(gc, cl)
    assert fl.n_points == len(pts)
    gp = np.array([[p[0], p[1]] for p in fl.points[:fl.n_points]], np.float32).reshape(-1, 2)
    assert np.allclose(gp, pts, atol=1e-3, rtol=0)  # device cos/sin vs libm: last-bit differences only


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["grid1", "lena", "blocky_720x1280", "ramps_480x640", "odd_33x129", "noise_123x77",
                                  "tiny_3x5", "row_1x64", "col_70x1"])
def test_cuda_hough_matches_oracle(ctx, name):
    img = all_cases()[name]
    thr = GOLD["cases"].get(name, {}).get("threshold", 5)
    for use_morph in (False, True):
        fl = ctx.hough(img, thr, use_morph=use_morph)[0]
        check_frame(fl, oracle_frame(img, thr, use_morph))


@pytest.mark.gpu
def test_cuda_hough_synthetic_grid_batch(ctx):
    K, D = K_for(640, 480), np.zeros(4)
    frames = np.stack([po.render_grid(R, t, K, D, 640, 480) for R, t in poses(6, seed=5)])
    out = ctx.hough(frames, 110)
    npts = 0
    for i in range(len(frames)):
        ref = oracle_frame(frames[i], 110)
        check_frame(out[i], ref)
        npts += len(ref[3])
    assert npts > 20


@pytest.mark.gpu
def test_cuda_hough_1080p_and_peak_overflow(ctx):
    img = all_cases()["blocky_1080x1920"]
    check_frame(ctx.hough(img, 200)[0], oracle_frame(img, 200))
    noisy = all_cases()["noise_480x640"]
    fl = ctx.hough(noisy, 2)[0]
    assert fl.flags == 1 and fl.n_peaks > 4096 and fl.n_lines == 1024
