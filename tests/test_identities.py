"""CPU: the two identities the CUDA path uses to skip work, checked in the ORACLE's arithmetic --
the reference's operations, restated -- so that they do not rest on the device code alone.

1. A shift hypothesis sees map point p where the unshifted pose sees p - shift: the WHITE score of (pose + shift) on
   the map equals the score of the pose on the map translated by -shift (search.cu: score_points / shift_sums).
2. Corner ordering 1 of a quad is the mirror of ordering 0: RPP gives R1 = R0 diag(1, -1, -1), t1 = t0, the same
   errors and the same number of iterations (pose.cu: k_pose_prep / k_pose_select)."""
import ctypes

import numpy as np

from frames import K_for, poses
from oracle import pyoracle as po


def _shifts(gs=0.32, n=9):
    half = n / 2.0 * gs
    out = []
    x = -half + gs / 2.0
    while x < half:          # accumulated exactly as computeAllShiftedHypothesesFAST writes it
        y = -half + gs / 2.0
        while y < half:
            out.append((x, y))
            y += gs
        x += gs
    return out


def test_shift_score_equals_score_on_translated_map():
    K, D = K_for(640, 480), np.zeros(4)
    white, red, green = po.make_map()
    shifts = _shifts()
    assert len(shifts) == 81
    for R, t in poses(2, seed=0x4D414E54 + 31):
        img = po.render_grid(R, t, K, D, 640, 480)
        tr, _ = po.localize(img, K, D)
        assert tr.status == 0
        best = np.concatenate([np.array(tr.optimized.w2c.R[:]), np.array(tr.optimized.w2c.t[:])])
        for k, (sx, sy) in enumerate(shifts):
            moved = best.copy()
            moved[9] += sx
            moved[10] += sy
            direct = po.score_white(moved, img, K, D)
            assert direct == tr.shift_err[k]               # what the pipeline itself scored for this shift
            d = np.array([sx, sy, 0.0])
            via_map = po.score_white(best, img, K, D, maps=(white - d, red - d, green - d))
            assert via_map == direct, (k, via_map, direct)


def test_second_corner_ordering_is_the_mirror_of_the_first():
    K, D = K_for(640, 480), np.zeros(4)
    s = 0.16
    model = [[(s, s), (-s, s), (-s, -s), (s, -s)], [(s, -s), (-s, -s), (-s, s), (s, s)]]   # gridSquarePossibilities
    mirror = np.diag([1.0, -1.0, -1.0])
    dbg_it = (ctypes.c_int * 8).in_dll(po.lib(), "mo_rpp_debug_iters")
    dbg_n = ctypes.c_int.in_dll(po.lib(), "mo_rpp_debug_n")
    n_quads = 0
    for R, t in poses(3, seed=0x4D414E54 + 32):
        tr, _ = po.localize(po.render_grid(R, t, K, D, 640, 480), K, D)
        for q in range(tr.n_quads):
            rays = list(tr.quad_test_norm[q])
            res = []
            for e in range(2):
                m, ip = np.zeros((3, 4)), np.ones((3, 4))
                for k in range(4):
                    m[0, k], m[1, k] = model[e][k]
                    ip[0, k], ip[1, k] = rays[2 * k], rays[2 * k + 1]
                ok, Rr, tt, _it, oe, ie = po.rpp(m, ip)
                res.append((ok, Rr, tt, oe, ie, [dbg_it[k] for k in range(dbg_n.value)]))
            (ok0, R0, t0, oe0, ie0, it0), (ok1, R1, t1, oe1, ie1, it1) = res
            assert ok0 == ok1
            if not ok0:
                continue
            n_quads += 1
            assert it0[0] == it1[0] and sorted(it0[1:]) == sorted(it1[1:])   # first run; candidate runs (roots come mirrored)
            assert np.abs(R1 - R0 @ mirror).max() < 1e-12 and np.abs(t1 - t0).max() < 1e-12
            assert abs(oe1 - oe0) < 1e-12 and abs(ie1 - ie0) < 1e-12
    assert n_quads > 100
