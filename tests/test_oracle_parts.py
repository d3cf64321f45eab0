"""CPU: the oracle's restated third-party pieces against cv2 4.13 (live, when importable), the
reference's own RPP (oracle/_ref/librpp_ref.so) and the reference's known-answer vectors."""
import ctypes
import os

import numpy as np
import pytest

from cases import all_cases
from frames import D_UNIT, K_UNIT, K_for, poses
from oracle import pyoracle as po

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_contours_and_approx_match_cv2():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(3)
    imgs = [po.edge_stage(all_cases()[k])[1] for k in ("grid1", "lena", "blocky_720x1280", "odd_33x129", "row_1x64")]
    for _ in range(25):
        h, w = rng.integers(1, 80), rng.integers(1, 80)
        imgs.append(((rng.random((h, w)) < rng.choice([0.2, 0.5, 0.7])) * 255).astype(np.uint8))
    imgs.append(((rng.random((200, 300)) < 0.55) * 255).astype(np.uint8))
    for img in imgs:
        for mcv, mo in ((cv2.RETR_CCOMP, po.RETR_CCOMP), (cv2.RETR_LIST, po.RETR_LIST)):
            cs, _ = cv2.findContours(img.copy(), mcv, cv2.CHAIN_APPROX_SIMPLE)
            oc, _ = po.find_contours(img, mo)
            assert len(cs) == len(oc)
            for a, b in zip(cs, oc):
                assert np.array_equal(a.reshape(-1, 2), b)  # same points, same order
        for c in cs:
            for eps in (10, 2.5):
                a = cv2.approxPolyDP(c, eps, True).reshape(-1, 2)
                b = po.approx_poly_dp(c.reshape(-1, 2), eps, True)
                assert np.array_equal(a, b)


def test_fisheye_matches_cv2():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(1)
    pts = np.concatenate([rng.random((500, 2)) * [1280, 1024], [[642.658203125, 501.5538330078125]]])
    for K, D in ((K_UNIT, D_UNIT), (K_for(640, 480), np.zeros(4))):
        und = cv2.fisheye.undistortPoints(pts.reshape(-1, 1, 2), K, D).reshape(-1, 2)
        assert np.array_equal(und, po.fisheye_undistort(pts, K, D))
        nrm = rng.normal(0, 1.0, (500, 2))
        dist = cv2.fisheye.distortPoints(nrm.reshape(-1, 1, 2), K, D).reshape(-1, 2)
        assert np.array_equal(dist, po.fisheye_distort(nrm, K, D))


def test_fisheye_roundtrip_unit_test_fixture():
    """test/unit_test.cpp:144-179: projecting through the fisheye model and undistorting is the identity."""
    nrm = np.array([[0.155, 0.855], [-0.155, 0.855], [-0.155, 0.545], [0.155, 0.545]])
    px = po.fisheye_distort(nrm, K_UNIT, D_UNIT)
    back = po.fisheye_undistort(px, K_UNIT, D_UNIT)
    assert np.abs(back - nrm).max() < 1e-7


def test_rng_matches_cv2():
    cv2 = pytest.importorskip("cv2")
    for seed in (1, 10, 12345):
        cv2.setRNGSeed(seed)
        a = np.empty((1, 4000), np.float32)
        cv2.randn(a, 0, 1)
        assert np.array_equal(a[0], po.rng_randn(seed, 4000))


def test_rng_golden_stream():
    """First draws of cv::RNG(1) / cv::RNG(10) (Mantis3Params.h:87, test/unit_test.cpp:196), recorded from cv2 4.13."""
    assert np.allclose(po.rng_randn(1, 3), [1.2680929e-10, 1.1373155, 2.2778807], rtol=1e-7, atol=0)
    assert np.allclose(po.rng_randn(10, 3), [3.1699796e-09, -1.2734357, -1.4336261], rtol=1e-7, atol=0)


# RobustPlanarPose/demo.cpp:17-38: the only real known-answer test in the reference
DEMO_MODEL = [0.0685, 0.6383, 0.4558, 0.7411, -0.7219, 0.7081, 0.7061, 0.2887, -0.9521, -0.2553,
              0.4636, 0.0159, -0.1010, 0.2817, 0.6638, 0.1582, 0.3925, -0.7954, 0.6965, -0.7795]
DEMO_IPRTS = [-0.0168, 0.0377, 0.0277, 0.0373, -0.0824, 0.0386, 0.0317, 0.0360, -0.1015, -0.0080,
              0.0866, 0.1179, 0.1233, 0.1035, 0.0667, 0.1102, 0.0969, 0.1660, 0.0622, 0.1608]
DEMO_R = np.array([[0.85763, -0.31179, 0.40898], [0.16047, -0.59331, -0.78882], [0.48859, 0.74214, -0.45881]])
DEMO_T = np.array([-0.10825, 1.26601, 11.19855])


def demo_inputs():
    model = np.zeros((3, 10))
    model[0], model[1] = DEMO_MODEL[:10], DEMO_MODEL[10:]
    ip = np.ones((3, 10))
    ip[0], ip[1] = DEMO_IPRTS[:10], DEMO_IPRTS[10:]
    return model, ip


def test_rpp_known_answer():
    ok, R, t, it, oe, ie = po.rpp(*demo_inputs())
    assert ok
    assert ((R - DEMO_R) ** 2).sum() < 1e-4 and ((t - DEMO_T) ** 2).sum() < 1e-4  # demo.cpp:107,114
    assert np.abs(R - DEMO_R).max() < 1e-5 and np.abs(t - DEMO_T).max() < 1e-5


def _ref_rpp():
    path = os.path.join(ROOT, "oracle", "_ref", "librpp_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/librpp_ref.so not built (reference tree absent)")
    L = ctypes.CDLL(path)

    def run(model, ip):
        m = np.ascontiguousarray(model, np.float64).copy()
        q = np.ascontiguousarray(ip, np.float64).copy()
        R, t = np.zeros(9), np.zeros(3)
        it, oe, ie = ctypes.c_int(), ctypes.c_double(), ctypes.c_double()
        ok = L.rpp_ref(m.ctypes.data, q.ctypes.data, m.shape[1], R.ctypes.data, t.ctypes.data, ctypes.byref(it),
                       ctypes.byref(oe), ctypes.byref(ie))
        return bool(ok), R.reshape(3, 3), t, it.value, oe.value, ie.value
    L.rpp_ref.argtypes = [ctypes.c_void_p] * 2 + [ctypes.c_int] + [ctypes.c_void_p] * 5
    return run


def test_rpp_matches_compiled_reference():
    ref = _ref_rpp()
    a, b = ref(*demo_inputs()), po.rpp(*demo_inputs())
    assert a[0] and b[0] and a[3] == b[3]
    assert np.abs(a[1] - b[1]).max() < 1e-12 and np.abs(a[2] - b[2]).max() < 1e-10
    rng = np.random.default_rng(0)
    s = 0.16
    squares = [np.array([[s, -s, -s, s], [s, s, -s, -s], [0, 0, 0, 0]]), np.array([[s, -s, -s, s], [-s, -s, s, s], [0, 0, 0, 0]])]
    close = total = 0
    for i, (Rw, tw) in enumerate(poses(400, seed=5)):
        R, t = Rw.T, -Rw.T @ tw  # world -> camera
        m = squares[i % 2]
        pc = R @ m + t[:, None]
        if (pc[2] <= 0.1).any():
            continue
        ip = pc / pc[2]
        ip[:2] += rng.normal(0, 0.003, (2, 4))
        a, b = ref(m, ip), po.rpp(m, ip)
        assert a[0] == b[0]
        if not a[0]:
            continue
        total += 1
        close += np.abs(a[1] - b[1]).max() < 1e-9 and np.abs(a[2] - b[2]).max() < 1e-9
        # both are fixed points of the same iteration: never further apart than its stop tolerance allows
        assert np.abs(a[1] - b[1]).max() < 5e-2 and abs(a[4] - b[4]) < 1e-5
    assert total > 300 and close >= 0.98 * total


def test_clean_mask_matches_cv2():
    cv2 = pytest.importorskip("cv2")
    for name in ("grid1", "lena", "odd_33x129", "noise_123x77", "ramps_480x640"):
        img = all_cases()[name]
        mono = cv2.GaussianBlur(cv2.cvtColor(img, cv2.COLOR_BGR2GRAY), (3, 3), 3, 3)
        canny = cv2.Canny(mono, 50, 150, apertureSize=3)
        tmp = cv2.morphologyEx(canny, cv2.MORPH_GRADIENT, cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (3, 3)))
        tmp = cv2.bitwise_not(tmp)
        small = np.zeros_like(tmp)
        cs, _ = cv2.findContours(tmp, cv2.RETR_LIST, cv2.CHAIN_APPROX_SIMPLE)
        for i in range(len(cs)):
            cv2.drawContours(small, cs, i, 255, 1)
        mask = cv2.bitwise_or(canny, small)
        for i in range(3):
            mask = cv2.dilate(mask, None, anchor=(-1, -1), iterations=3 + i, borderType=1, borderValue=1)
            mask = cv2.erode(mask, None, anchor=(-1, -1), iterations=3 + i, borderType=1, borderValue=1)
        mask = cv2.erode(mask, None, anchor=(-1, -1), iterations=3, borderType=1, borderValue=1)
        assert np.array_equal(mask, po.clean_mask(canny)), name


def test_map_matches_reference_yaml():
    path = "/root/reference/params/map.yaml"
    if not os.path.exists(path):
        pytest.skip("reference tree absent")
    import re

    s = open(path).read()
    w, r, g = po.make_map()
    for name, arr in (("whiteMap", w), ("redMap", r), ("greenMap", g)):
        rows = [x.replace("\n", "").replace(" ", "") for x in re.search(name + r": '([^']*)'", s).group(1).split(";")]
        ref = np.array([[float(v) for v in x.split(",")] for x in rows if x])
        assert np.array_equal(ref, arr)
    assert (len(w), len(r), len(g)) == (646, 37, 37)
