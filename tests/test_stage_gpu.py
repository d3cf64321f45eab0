"""GPU: the stage-level entry points of the pose and scoring stages through the C ABI:
mantis_rpp_batch against the reference's own KAT (RobustPlanarPose/demo.cpp:17-38, threshold 1e-4), against the
reference's RPP.cpp compiled into oracle/_ref/librpp_ref.so, and against the oracle (R, t, img_err, iteration
counts); mantis_score_batch against the oracle's evaluateHypothesisWithImageWHITE (exact: integer sums)."""
import numpy as np
import pytest

from frames import D_UNIT, K_for, poses
from oracle import pyoracle as po
from test_oracle_parts import DEMO_R, DEMO_T, _ref_rpp, demo_inputs

pytestmark = pytest.mark.gpu


def _abi(model3xn, ip3xn):
    """3 x n row-major (the reference's cv::Mat layout) -> the ABI's point-major arrays."""
    return np.ascontiguousarray(model3xn.T), np.ascontiguousarray(ip3xn[:2].T)


def test_rpp_reference_known_answer(ctx):
    """demo.cpp: 10 planar points, ground truth from the original Matlab code, pass threshold 1e-4 squared error."""
    model, ip = demo_inputs()
    m, q = _abi(model, ip)
    r = ctx.rpp(m[None], q[None])[0]
    assert r.status == 0
    R, t = np.array(r.R[:]).reshape(3, 3), np.array(r.t[:])
    assert ((R - DEMO_R) ** 2).sum() < 1e-4 and ((t - DEMO_T) ** 2).sum() < 1e-4   # demo.cpp:107,114
    ok, Ro, to, it, oe, ie = po.rpp(model, ip)
    assert np.abs(R - Ro).max() < 1e-9 and np.abs(t - to).max() < 1e-8
    assert r.obj_err == pytest.approx(oe, rel=1e-6) and r.img_err == pytest.approx(ie, rel=1e-9)
    ref = _ref_rpp()(model, ip)
    assert np.abs(R - ref[1]).max() < 1e-9 and np.abs(t - ref[2]).max() < 1e-8


def _squares(n, seed, noise=0.003):
    rng = np.random.default_rng(seed)
    s = 0.16
    sq = [np.array([[s, -s, -s, s], [s, s, -s, -s], [0, 0, 0, 0.0]]), np.array([[s, -s, -s, s], [-s, -s, s, s], [0, 0, 0, 0.0]])]
    out = []
    for i, (Rw, tw) in enumerate(poses(n, seed=seed)):
        R, t = Rw.T, -Rw.T @ tw
        m = sq[i % 2]
        pc = R @ m + t[:, None]
        if (pc[2] <= 0.1).any():
            continue
        ip = pc / pc[2]
        ip[:2] += rng.normal(0, noise, (2, 4))
        out.append((m, ip))
    return out


def test_rpp_400_squares_against_compiled_reference_and_oracle(ctx):
    """Planar 4-point problems = the quads of the localization path, through its own kernels (k_pose_run /
    k_pose_second): R, t, img_err and the iteration counts of both ObjPose stages, per problem."""
    ref = _ref_rpp()
    cases = _squares(400, 5)
    mm = np.stack([_abi(m, ip)[0] for m, ip in cases])
    qq = np.stack([_abi(m, ip)[1] for m, ip in cases])
    out = ctx.rpp(mm, qq)
    olib = po.lib()
    import ctypes
    dbg_it = (ctypes.c_int * 8).in_dll(olib, "mo_rpp_debug_iters")
    dbg_n = ctypes.c_int.in_dll(olib, "mo_rpp_debug_n")
    n_ok = ref_close = 0
    for k, (m, ip) in enumerate(cases):
        r = out[k]
        ok, Ro, to, it, oe, ie = po.rpp(m, ip)
        o_its = [dbg_it[j] for j in range(dbg_n.value)]
        assert (r.status == 0) == ok, k
        if not ok:
            continue
        n_ok += 1
        R, t = np.array(r.R[:]).reshape(3, 3), np.array(r.t[:])
        # the oracle: same fixed point, same number of steps
        assert np.abs(R - Ro).max() < 1e-9 and np.abs(t - to).max() < 1e-9, k
        assert r.img_err == pytest.approx(ie, rel=1e-7, abs=1e-12) and r.obj_err == pytest.approx(oe, rel=1e-5, abs=1e-12)
        assert r.iters_first == o_its[0] and r.n_candidates == len(o_its) - 1, (k, r.iters_first, o_its)
        assert r.iters_best in o_its[1:]
        # the compiled reference (Jenkins-Traub roots, cv::SVD): the same minimiser unless its stop test fires elsewhere
        a = ref(m, ip)
        assert a[0]
        assert np.abs(R - a[1]).max() < 5e-2 and abs(r.obj_err - a[4]) < 1e-5
        ref_close += np.abs(R - a[1]).max() < 1e-9 and np.abs(t - a[2]).max() < 1e-9
    assert n_ok > 300 and ref_close >= 0.98 * n_ok


def test_rpp_general_shapes(ctx):
    """3 .. 16 points, off-plane models: the general (3x3 SVD) kernel against the oracle."""
    rng = np.random.default_rng(17)
    for m_pts in (5, 8, 16):
        models, ips = [], []
        for Rw, tw in poses(12, seed=100 + m_pts):
            R, t = Rw.T, -Rw.T @ tw
            m = np.vstack([rng.uniform(-0.4, 0.4, (2, m_pts)), np.zeros((1, m_pts))])
            pc = R @ m + t[:, None]
            ip = pc / pc[2]
            ip[:2] += rng.normal(0, 0.001, (2, m_pts))
            models.append(m); ips.append(ip)
        out = ctx.rpp(np.stack([_abi(a, b)[0] for a, b in zip(models, ips)]), np.stack([_abi(a, b)[1] for a, b in zip(models, ips)]))
        for k in range(len(models)):
            ok, Ro, to, it, oe, ie = po.rpp(models[k], ips[k])
            assert (out[k].status == 0) == ok
            if ok:
                assert np.abs(np.array(out[k].R[:]).reshape(3, 3) - Ro).max() < 1e-8
                assert np.abs(np.array(out[k].t[:]) - to).max() < 1e-8


@pytest.mark.parametrize("dist", [False, True])
def test_score_batch_matches_oracle(ctx, dist):
    """Scores of given (pose, image) pairs: exact (integer colour sums over the same pixels)."""
    K = K_for(640, 480)
    D = D_UNIT if dist else np.zeros(4)
    ps = poses(3, seed=0x4D414E54 + 41)
    frames = np.stack([po.render_grid(R, t, K, D, 640, 480) for R, t in ps])
    rng = np.random.default_rng(3)
    n_h = 130   # more than one pass of 96
    w2c = np.zeros((len(ps), n_h, 12))
    for i, (R, t) in enumerate(ps):
        for h in range(n_h):
            a = rng.normal(0, 0.02, 3)
            dR = po.rodrigues_exp(a)
            w2c[i, h, :9] = (R @ dR).reshape(-1)
            w2c[i, h, 9:] = t + rng.normal(0, 0.02, 3)
        w2c[i, -1, 9:] = [50.0, 50.0, 1.0]   # looks at nothing: DBL_MAX or a black-floor score, as the oracle says
    got = ctx.score(frames, K, D, w2c)
    for i in range(len(ps)):
        for h in list(range(0, n_h, 7)) + [n_h - 1]:
            assert got[i, h] == po.score_white(w2c[i, h], frames[i], K, D), (i, h)
