"""the 2016 generation's particle weight + undistortImage, and the Markov yaw filter.
CPU: the oracle against cv2.fisheye.undistortImage (bit-exact) and closed-form properties of the filter.
GPU: the CUDA entry points against the oracle."""
import numpy as np
import pytest

from cases import all_cases
from frames import D_UNIT, K_for, poses
from oracle import pyoracle as po

K_LENA = np.array([[200, 0, 256], [0, 200, 256], [0, 0, 1]], np.float32)


def test_oracle_undistort_matches_cv2_bit_exact():
    cv2 = pytest.importorskip("cv2")
    for name, K in (("lena", K_LENA), ("grid1", np.array([[229.0, 0, 453.5], [0, 228.75, 252.5], [0, 0, 1]], np.float32))):
        img = all_cases()[name]
        ref = cv2.fisheye.undistortImage(img, K.astype(np.float64), D_UNIT, Knew=K.astype(np.float64))
        out, m1, m2 = po.undistort_image(img, K, D_UNIT)
        assert np.array_equal(ref, out)
        r1, r2 = cv2.fisheye.initUndistortRectifyMap(K.astype(np.float64), D_UNIT, np.eye(3), K.astype(np.float64),
                                                     (img.shape[1], img.shape[0]), cv2.CV_16SC2)
        assert np.array_equal(r1, m1) and np.array_equal(r2, m2)


def test_oracle_markov_properties():
    m = po.Markov(yaw_rad=np.deg2rad(120.4))
    assert abs(m.p.sum() - 1) < 1e-12 and m.get_yaw()[1] == 120
    # Gaussian of sigma 3 around the bin: ratio of neighbours = exp(-1/18)
    assert m.p[121] / m.p[120] == pytest.approx(np.exp(-1 / 18), rel=1e-9)
    m.convolve(np.deg2rad(45), 5.0)          # (int)dTheta == 0: no shift, only diffusion (the reference's cast)
    assert m.get_yaw()[1] == 120 and abs(m.p.sum() - 1) < 1e-12
    m.convolve(1.2, 1.0)                     # (int)1.2 rad = 1 -> 57 bins
    assert m.get_yaw()[1] == 177
    m.sense(np.deg2rad(-170))                # bin 190
    assert 177 < m.get_yaw()[1] <= 190
    e = m.update_errors([np.deg2rad(185.5), np.deg2rad(10)], [2.0, 2.0])
    assert e[0] == 2.0 / m.p[185] and e[1] > e[0]


def _ref_markov():
    import ctypes
    import os

    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libmarkov_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libmarkov_ref.so not built (reference tree absent)")
    L = ctypes.CDLL(path)
    d, vp, ci = ctypes.c_double, ctypes.c_void_p, ctypes.c_int
    L.markov_ref_update_weights.argtypes = [vp, d]
    L.markov_ref_init.argtypes = [vp, d]
    L.markov_ref_convolve.argtypes = [vp, d, d]
    L.markov_ref_sense_plane.argtypes = [vp, vp]
    L.markov_ref_update_errors.argtypes = [vp, vp, vp, ci]
    L.markov_ref_get_yaw.argtypes = [vp]
    L.markov_ref_get_yaw.restype = d
    return L


def test_oracle_markov_matches_compiled_reference():
    """oracle/legacy.c against the reference's own include/mantis3/Markov.cpp, compiled unmodified into
    oracle/_ref/libmarkov_ref.so (oracle/ref_build): constructor, updateWeights, convolve, senseFusion(plane),
    updateHypothesis and getYaw on seeded sequences -- the same doubles, bit for bit."""
    L = _ref_markov()
    rng = np.random.default_rng(0x4D414E54 + 71)
    for trial in range(12):
        yaw0 = rng.uniform(-np.pi, np.pi)
        m = po.Markov(yaw_rad=yaw0)
        r = np.zeros(360)
        L.markov_ref_init(r.ctypes.data, float(yaw0))
        assert np.array_equal(m.p, r), ("init", trial)
        for step in range(6):
            dth, dt = rng.uniform(-2.5, 2.5), rng.uniform(0.5, 8.0)
            m.convolve(dth, dt)
            L.markov_ref_convolve(r.ctypes.data, float(dth), float(dt))
            assert np.array_equal(m.p, r), ("convolve", trial, step)
            # senseFusion(Hypothesis) cannot run in the reference (its zeroing loop never ends, Markov.cpp:190-191):
            # the sense plane is built as that function means to, with the reference's own updateWeights
            ys = rng.uniform(-np.pi, np.pi)
            sense = np.zeros(360)
            deg = ys * 180.0 / np.pi
            sense[int(deg + 360 if deg < 0 else deg)] = 1
            L.markov_ref_update_weights(sense.ctypes.data, 3.5)
            L.markov_ref_sense_plane(r.ctypes.data, sense.ctypes.data)
            m.sense(ys)
            assert np.array_equal(m.p, r), ("sense", trial, step)
            yaws = rng.uniform(-np.pi, np.pi, 9)
            errs = rng.uniform(1e3, 1e5, 9)
            e_ref = errs.copy()
            L.markov_ref_update_errors(r.ctypes.data, yaws.ctypes.data, e_ref.ctypes.data, 9)
            assert np.array_equal(m.update_errors(yaws, errs), e_ref), ("update_errors", trial, step)
            assert m.get_yaw()[0] == L.markov_ref_get_yaw(r.ctypes.data)
    # updateWeights on an arbitrary plane, several widths
    for sd in (0.4, 3.0, 3.5, 11.0):
        plane = rng.uniform(0, 1, 360)
        a = po.Markov(plane=plane)
        a.update_weights(sd)
        b = plane.copy()
        L.markov_ref_update_weights(b.ctypes.data, float(sd))
        assert np.array_equal(a.p, b), sd


def _particles(n_p, seed):
    rng = np.random.default_rng(seed)
    out = []
    for R, t in poses(n_p, seed=seed):
        Rc, tc = R.T, -R.T @ t           # world -> camera
        out.append(np.concatenate([Rc.reshape(-1), tc]))
    return np.array(out) + rng.normal(0, 1e-3, (n_p, 12))


def test_oracle_camera_error_counts_and_floor():
    K = K_for(640, 480)
    w, r, g = po.make_map()
    img = np.full((480, 640, 3), 255, np.uint8)
    T = _particles(1, 5)[0]
    e = po.camera_error(img, K, T, w, r, g)
    assert e >= 0 and e < 3 * 255 ** 2
    behind = T.copy(); behind[9:] = [0, 0, -50]          # nothing in frame -> 1e17 / 0
    far = po.camera_error(img, K, np.concatenate([np.eye(3).reshape(-1), [100, 100, 1]]), w, r, g)
    assert far == np.inf or far >= 1e15


def _ref_montecarlo():
    import ctypes
    import os
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libmontecarlo_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libmontecarlo_ref.so not built (reference tree absent)")
    L = ctypes.CDLL(path)
    vp, ci, d = ctypes.c_void_p, ctypes.c_int, ctypes.c_double
    L.montecarlo_ref_camera_error.argtypes = [vp, ci, ci, ci, vp, vp, vp, ci, ci, ci]
    L.montecarlo_ref_camera_error.restype = d
    L.montecarlo_ref_weight.argtypes = [vp, ci, ci, ci, vp, vp, vp, vp, ci, ci, ci]
    L.montecarlo_ref_weight.restype = d
    return L


def test_oracle_camera_error_matches_compiled_reference():
    """oracle/legacy.c's particle weight against the reference's own MonteCarlo.cpp (project2d, colorError,
    computeCameraError, computeWeight), compiled unmodified into oracle/_ref/libmontecarlo_ref.so (oracle/ref_build,
    OpenCV / ROS / tf stand-ins in shim_mc): rendered frames and noise, poses around the frame's own and far off it,
    a projection that rounds to column `cols` / row `rows`, and the fewer-than-10-points floor.  Equal doubles."""
    L = _ref_montecarlo()
    K, D = K_for(640, 480), np.zeros(4)
    Kf = np.ascontiguousarray(K, np.float32)
    w, r, g = po.make_map()
    mp = np.ascontiguousarray(np.concatenate([w, r, g]), np.float64)
    rng = np.random.default_rng(17)
    frames = [po.render_grid(R, t, K, D, 640, 480) for R, t in poses(2, seed=31)]
    frames.append(rng.integers(0, 256, (480, 640, 3), dtype=np.uint8))
    n_checked = 0
    for i, img in enumerate(frames):
        img = np.ascontiguousarray(img)
        P = list(_particles(60, 200 + i))
        P.append(np.concatenate([np.eye(3).reshape(-1), [100, 100, 1]]))      # nothing projected: 1e17 / 0
        P.append(np.concatenate([np.eye(3).reshape(-1), [0.0, 0.0, 1.0]]))    # a few points in frame
        for T in P:
            T = np.ascontiguousarray(T, np.float64)
            ref = L.montecarlo_ref_camera_error(img.ctypes.data, 640, 480, 3 * 640, Kf.ctypes.data, T.ctypes.data, mp.ctypes.data,
                                                len(w), len(r), len(g))
            got = po.camera_error(img, K, T, w, r, g)
            assert got == ref or (np.isinf(ref) and np.isinf(got)) or (np.isnan(ref) and np.isnan(got)), (i, got, ref)
            n_checked += 1
    # one map point whose projection rounds to x == cols and one to y == rows (Mat::at is unchecked)
    one = np.array([[0.0, 0.0, 0.0]] * 12)
    img = np.ascontiguousarray(frames[2])
    for tx, ty, black in ((319.8 / 320.0, 0.0, False), (0.0, 239.7 / 320.0, True)):
        T = np.concatenate([np.eye(3).reshape(-1), [tx, ty, 1.0]])
        ref = L.montecarlo_ref_camera_error(img.ctypes.data, 640, 480, 3 * 640, Kf.ctypes.data, T.ctypes.data, one.ctypes.data, 12, 0, 0)
        got = po.camera_error(img, K, T, one, one[:0], one[:0])
        assert got == ref
        # x == cols is the first pixel of the next row (linear addressing); y == rows lies below the image: black
        expect = 3 * 255 ** 2 if black else int(((img[241, 0].astype(int) - 255) ** 2).sum())
        assert got == float(expect)
    # computeWeight: particle (world <- base) and base -> camera
    for k, (R, t) in enumerate(poses(8, seed=77)):
        b2c = np.concatenate([np.array([[0, -1, 0], [-1, 0, 0], [0, 0, -1.0]]).reshape(-1), [0.02, -0.01, 0.05]])
        part = np.concatenate([R.reshape(-1), t])
        ref = L.montecarlo_ref_weight(img.ctypes.data, 640, 480, 3 * 640, Kf.ctypes.data, part.ctypes.data, b2c.ctypes.data,
                                      mp.ctypes.data, len(w), len(r), len(g))
        Rp, tp, Rb, tb = R, t, b2c[:9].reshape(3, 3), b2c[9:]
        # inverse() and the product as tf composes them: inv = (Rp^T, -Rp^T tp), b2c_inv = (Rb^T, -Rb^T tb)
        Ri, ti = Rp.T, -(Rp.T @ tp)
        Rbi, tbi = Rb.T, -(Rb.T @ tb)
        T = np.concatenate([(Ri @ Rbi).reshape(-1), Ri @ tbi + ti])
        got = po.camera_error(img, K, T, w, r, g)
        assert got == pytest.approx(ref, rel=1e-9) or (np.isinf(ref) and np.isinf(got)), (k, got, ref)
    assert n_checked == 3 * 62


@pytest.mark.gpu
def test_cuda_undistort_matches_oracle(ctx):
    for name, K in (("lena", K_LENA), ("odd_33x129", K_for(129, 33))):
        img = all_cases()[name]
        got = ctx.undistort(np.stack([img, img[::-1].copy()]), K, D_UNIT)
        assert np.array_equal(got[0], po.undistort_image(img, K, D_UNIT)[0])
        assert np.array_equal(got[1], po.undistort_image(img[::-1].copy(), K, D_UNIT)[0])


@pytest.mark.gpu
def test_cuda_particle_weights_match_oracle(ctx):
    K, D = K_for(640, 480), np.zeros(4)
    w, r, g = po.make_map()
    frames = np.stack([po.render_grid(R, t, K, D, 640, 480) for R, t in poses(3, seed=31)])
    P = np.stack([_particles(40, 100 + i) for i in range(3)])
    P[0, 0] = np.concatenate([np.eye(3).reshape(-1), [100, 100, 1]])   # nothing projected
    got = ctx.particle_weights(frames, K, P)
    for i in range(3):
        for k in range(40):
            ref = po.camera_error(frames[i], K, P[i, k], w, r, g)
            assert got[i, k] == ref or (np.isinf(ref) and np.isinf(got[i, k])), (i, k, got[i, k], ref)


@pytest.mark.gpu
def test_cuda_markov_matches_oracle(ctx):
    yaws = np.deg2rad([120.4, -30.0, 359.2, 0.0, 181.0])
    planes = ctx.markov(np.zeros((5, 360)), ctx.MARKOV_INIT, yaws)
    ref = [po.Markov(yaw_rad=y) for y in yaws]
    dth, dt = np.array([0.2, 1.2, -2.5, 3.1, -0.4]), np.array([0.5, 1.0, 2.0, 0.1, 5.0])
    planes = ctx.markov(planes, ctx.MARKOV_CONVOLVE, dth, dt)
    sense = np.deg2rad([118.0, 40.0, -5.0, 170.0, -179.0])
    planes2 = ctx.markov(planes, ctx.MARKOV_SENSE, sense)
    for i, m in enumerate(ref):
        m.convolve(dth[i], dt[i])
        assert np.allclose(planes[i], m.p, rtol=1e-10, atol=1e-300)
        m.sense(sense[i])
        assert np.allclose(planes2[i], m.p, rtol=1e-9, atol=1e-300)
        assert abs(planes2[i].sum() - 1) < 1e-9
    so = np.array([0, 0, 1, 4, 3], np.int32)
    hy = np.deg2rad([119.0, 10.0, 41.0, -179.5, 171.0])
    err = ctx.markov_update_errors(planes2, so, hy, np.full(5, 3.0))
    for k in range(5):
        assert err[k] == pytest.approx(ref[so[k]].update_errors([hy[k]], [3.0])[0], rel=1e-9)
