"""ctypes binding of the C oracle (oracle/libmantis_oracle.so).

TEST INFRASTRUCTURE ONLY -- see oracle/mantis_oracle.h.  The product path
(mantis_b200/) never imports this module.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force=False):
    so = os.path.join(_HERE, "libmantis_oracle.so")
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".c", ".h"))]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(build())
    return _LIB


def _u8(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    return a, a.ctypes.data_as(ctypes.c_void_p)


def bgr2gray(bgr):
    bgr, p = _u8(bgr)
    h, w, _ = bgr.shape
    out = np.empty((h, w), np.uint8)
    lib().mo_bgr2gray(p, w, h, w * 3, out.ctypes.data_as(ctypes.c_void_p))
    return out


def blur3(gray):
    gray, p = _u8(gray)
    h, w = gray.shape
    out = np.empty((h, w), np.uint8)
    lib().mo_blur3(p, w, h, out.ctypes.data_as(ctypes.c_void_p))
    return out


def sobel3(gray):
    gray, p = _u8(gray)
    h, w = gray.shape
    dx = np.empty((h, w), np.int16)
    dy = np.empty((h, w), np.int16)
    lib().mo_sobel3(p, w, h, dx.ctypes.data_as(ctypes.c_void_p), dy.ctypes.data_as(ctypes.c_void_p))
    return dx, dy


def canny(gray, low, high):
    gray, p = _u8(gray)
    h, w = gray.shape
    out = np.empty((h, w), np.uint8)
    lib().mo_canny(p, w, h, int(low), int(high), out.ctypes.data_as(ctypes.c_void_p))
    return out


def dilate3(img, iters):
    img, p = _u8(img)
    h, w = img.shape
    out = np.empty((h, w), np.uint8)
    lib().mo_dilate3(p, w, h, int(iters), out.ctypes.data_as(ctypes.c_void_p))
    return out


def erode3(img, iters):
    img, p = _u8(img)
    h, w = img.shape
    out = np.empty((h, w), np.uint8)
    lib().mo_erode3(p, w, h, int(iters), out.ctypes.data_as(ctypes.c_void_p))
    return out


def edge_stage(bgr, canny_low=50):
    """gray -> blur -> Canny(T,3T) -> dilate x2 -> erode x1.  Returns (canny, morph)."""
    bgr, p = _u8(bgr)
    h, w, _ = bgr.shape
    can = np.empty((h, w), np.uint8)
    mor = np.empty((h, w), np.uint8)
    lib().mo_edge_stage(p, w, h, w * 3, int(canny_low), can.ctypes.data_as(ctypes.c_void_p),
                        mor.ctypes.data_as(ctypes.c_void_p))
    return can, mor


# ---- contours ----
class _Contour(ctypes.Structure):
    _fields_ = [("first", ctypes.c_int), ("count", ctypes.c_int), ("is_hole", ctypes.c_int),
                ("parent", ctypes.c_int), ("discovery", ctypes.c_int)]


class _Contours(ctypes.Structure):
    _fields_ = [("n", ctypes.c_int), ("npts", ctypes.c_int), ("c", ctypes.POINTER(_Contour)),
                ("pts", ctypes.POINTER(ctypes.c_int))]


RETR_LIST, RETR_CCOMP = 1, 2


def find_contours(img, mode=RETR_CCOMP):
    """Returns (list of int32 [k,2] point arrays, list of is_hole flags) in cv2's output order."""
    img, p = _u8(img)
    h, w = img.shape
    out = _Contours()
    lib().mo_find_contours(p, w, h, int(mode), ctypes.byref(out))
    pts = np.ctypeslib.as_array(out.pts, shape=(max(out.npts, 1), 2)).copy()
    res, holes = [], []
    for i in range(out.n):
        c = out.c[i]
        res.append(pts[c.first:c.first + c.count].astype(np.int32))
        holes.append(bool(c.is_hole))
    lib().mo_free_contours(ctypes.byref(out))
    return res, holes


def approx_poly_dp(pts, eps, closed=True):
    pts = np.ascontiguousarray(pts, dtype=np.int32).reshape(-1, 2)
    dst = np.empty_like(pts)
    f = lib().mo_approx_poly_dp
    f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_double, ctypes.c_int, ctypes.c_void_p]
    n = f(pts.ctypes.data, len(pts), float(eps), int(closed), dst.ctypes.data)
    return dst[:n].copy()


# ---- camera / rng ----
def _dp(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def fisheye_undistort(pts, K, D):
    pts = np.ascontiguousarray(pts, np.float64).reshape(-1, 2)
    K = np.ascontiguousarray(K, np.float32)
    D = np.ascontiguousarray(D, np.float64)
    out = np.empty_like(pts)
    lib().mo_fisheye_undistort(_dp(pts), len(pts), _dp(K), _dp(D), _dp(out))
    return out


def fisheye_distort(pts, K, D):
    pts = np.ascontiguousarray(pts, np.float64).reshape(-1, 2)
    K = np.ascontiguousarray(K, np.float32)
    D = np.ascontiguousarray(D, np.float64)
    out = np.empty_like(pts)
    lib().mo_fisheye_distort(_dp(pts), len(pts), _dp(K), _dp(D), _dp(out))
    return out


def rng_randn(seed, n):
    f = lib().mo_rng_randn
    f.restype = ctypes.c_float
    st = ctypes.c_uint64(seed if seed else 0xFFFFFFFF)
    return np.array([f(ctypes.byref(st)) for _ in range(n)], np.float32)


def ziggurat_tables():
    kn = np.empty(128, np.uint32)
    wn = np.empty(128, np.float32)
    fn = np.empty(128, np.float32)
    lib().mo_ziggurat_tables(_dp(kn), _dp(wn), _dp(fn))
    return kn, wn, fn


def rpp(model3xn, iprts3xn):
    """Returns (ok, R[3,3], t[3], iters, obj_err, img_err)."""
    m = np.ascontiguousarray(model3xn, np.float64)
    ip = np.ascontiguousarray(iprts3xn, np.float64)
    R = np.zeros(9)
    t = np.zeros(3)
    it = ctypes.c_int()
    oe = ctypes.c_double()
    ie = ctypes.c_double()
    ok = lib().mo_rpp(_dp(m), _dp(ip), m.shape[1], _dp(R), _dp(t), ctypes.byref(it), ctypes.byref(oe), ctypes.byref(ie))
    return bool(ok), R.reshape(3, 3), t, it.value, oe.value, ie.value


# ---- map (params/map.yaml geometry; params/matGen_dense.m) ----
def make_map(spacing=0.32):
    """white [646,3], red [37,3], green [37,3]: the values params/map.yaml holds (printed with
    %f by matGen_dense.m, parsed back with atof by parseCoordinatesFromString)."""
    r6 = lambda v: float("%f" % v)
    white = []
    for y in np.arange(-3.5, 3.5 + 1e-9, 1.0):
        for x in np.arange(-4.5, 4.5 + 1e-9, 0.25):
            white.append((r6(spacing * x), r6(spacing * y), 0.0))
    for x in np.arange(-4.5, 4.5 + 1e-9, 1.0):
        for y in np.arange(-4.25, 4.25 + 1e-9, 0.25):
            white.append((r6(spacing * x), r6(spacing * y), 0.0))
    green = [(r6(spacing * x), r6(4.5 * spacing), 0.0) for x in np.arange(-4.5, 4.5 + 1e-9, 0.25)]
    red = [(r6(spacing * x), r6(-4.5 * spacing), 0.0) for x in np.arange(-4.5, 4.5 + 1e-9, 0.25)]
    return np.array(white), np.array(red), np.array(green)


# ---- pipeline ----
MAX_QUADS = 1024


class _Tf(ctypes.Structure):
    _fields_ = [("R", ctypes.c_double * 9), ("t", ctypes.c_double * 3)]


class _Hyp(ctypes.Structure):
    _fields_ = [("c2w", _Tf), ("w2c", _Tf), ("q", ctypes.c_double * 4), ("error", ctypes.c_double)]


class Params(ctypes.Structure):
    _fields_ = [("canny_hysteresis", ctypes.c_int), ("polygon_epsilon", ctypes.c_int),
                ("search_radius_multiplier", ctypes.c_double), ("grid_spacing", ctypes.c_double),
                ("grid_size", ctypes.c_int), ("quad_stretch_factor", ctypes.c_double),
                ("max_quad_error", ctypes.c_double), ("max_angle_difference", ctypes.c_double),
                ("projection_bias", ctypes.c_double), ("color_projection_bias", ctypes.c_double),
                ("rot_sigma", ctypes.c_double), ("trans_sigma", ctypes.c_double),
                ("particles", ctypes.c_int), ("iterations", ctypes.c_int), ("rng_seed", ctypes.c_uint64)]


class _Map(ctypes.Structure):
    _fields_ = [("n_white", ctypes.c_int), ("n_red", ctypes.c_int), ("n_green", ctypes.c_int),
                ("white", ctypes.c_void_p), ("red", ctypes.c_void_p), ("green", ctypes.c_void_p)]


class Trace(ctypes.Structure):
    _fields_ = [("status", ctypes.c_int), ("n_contours", ctypes.c_int), ("n_quads_raw", ctypes.c_int),
                ("n_quads", ctypes.c_int),
                ("quad_corners", (ctypes.c_int * 8) * MAX_QUADS),
                ("quad_test_px", (ctypes.c_double * 8) * MAX_QUADS),
                ("quad_test_norm", (ctypes.c_double * 8) * MAX_QUADS),
                ("n_hyps_generated", ctypes.c_int), ("n_hyps_cluster", ctypes.c_int),
                ("quad_pass", ctypes.c_int * MAX_QUADS), ("quad_img_err", ctypes.c_double * MAX_QUADS),
                ("first_best", _Hyp), ("optimized", _Hyp), ("shift_err", ctypes.c_double * 81),
                ("yaw_err", ctypes.c_double * 4), ("best_rot", ctypes.c_int), ("min_yaw_diff", ctypes.c_double),
                ("final_hyp", _Hyp), ("pos", ctypes.c_double * 3), ("quat_xyzw", ctypes.c_double * 4),
                ("error", ctypes.c_double), ("n_scored", ctypes.c_long)]


def default_params():
    p = Params()
    lib().mo_default_params(ctypes.byref(p))
    return p


_MAP_KEEP = {}


def _map_struct(maps=None):
    key = id(maps)
    if key not in _MAP_KEEP:
        w, r, g = maps if maps is not None else make_map()
        w, r, g = (np.ascontiguousarray(a, np.float64) for a in (w, r, g))
        m = _Map(len(w), len(r), len(g), w.ctypes.data, r.ctypes.data, g.ctypes.data)
        _MAP_KEEP[key] = (m, w, r, g, maps)   # (holding `maps` keeps its id from being handed to another tuple)
    return _MAP_KEEP[key][0]


def localize(bgr, K, D, params=None, maps=None, want_mask=False):
    """Whole-frame oracle: returns (Trace, mask or None)."""
    bgr, p = _u8(bgr)
    h, w, _ = bgr.shape
    K = np.ascontiguousarray(K, np.float32)
    D = np.ascontiguousarray(D, np.float64)
    prm = params if params is not None else default_params()
    tr = Trace()
    mask = np.empty((h, w), np.uint8) if want_mask else None
    lib().mo_localize(p, w, h, 3 * w, _dp(K), _dp(D), ctypes.byref(prm), ctypes.byref(_map_struct(maps)),
                      ctypes.byref(tr), _dp(mask) if want_mask else None)
    return tr, mask


def localize_rig(frames, Ks, Ds, cam_to_base, params=None, maps=None):
    """Joint localization of one rig: frames [C,H,W,3], Ks [C,3,3], Ds [C,4], cam_to_base [C,7] (x y z qx qy qz qw of
    each camera in the base frame; None = a single camera without extrinsics).  Returns the Trace (base pose)."""
    frames = np.ascontiguousarray(frames, np.uint8)
    C, h, w, _ = frames.shape
    Ks = np.ascontiguousarray(Ks, np.float32).reshape(C, 9)
    Ds = np.ascontiguousarray(Ds, np.float64).reshape(C, 4)
    ptrs = (ctypes.c_void_p * C)(*[frames[k].ctypes.data for k in range(C)])
    b2c = None if cam_to_base is None else np.ascontiguousarray(cam_to_base, np.float64).reshape(C, 7)
    prm = params if params is not None else default_params()
    tr = Trace()
    lib().mo_localize_rig(ptrs, C, w, h, 3 * w, _dp(Ks), _dp(Ds), _dp(b2c) if b2c is not None else None, ctypes.byref(prm),
                          ctypes.byref(_map_struct(maps)), ctypes.byref(tr))
    return tr


def score_white(w2c12, bgr, K, D, maps=None, bias=1.1):
    """evaluateHypothesisWithImageWHITE on cleanImageByEdge's image (HypothesisEvaluation.h:107-158, 319-386) for the
    hypothesis whose W2C is w2c12 (R row-major + t)."""
    bgr, p = _u8(bgr)
    h, w, _ = bgr.shape
    K = np.ascontiguousarray(K, np.float32)
    D = np.ascontiguousarray(D, np.float64)
    can, _m = edge_stage(bgr, 50)
    mask = clean_mask(can)
    tf = _Tf()
    v = np.asarray(w2c12, np.float64).reshape(-1)
    tf.R[:] = v[:9].tolist()
    tf.t[:] = v[9:].tolist()
    hyp = _Hyp()
    lib().mo_hyp_set_w2c(ctypes.byref(hyp), ctypes.byref(tf))
    lib().mo_score_white.restype = ctypes.c_double
    lib().mo_score_white.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                     ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_double]
    return lib().mo_score_white(ctypes.byref(hyp), p, mask.ctypes.data, w, h, 3 * w, K.ctypes.data, D.ctypes.data,
                                ctypes.addressof(_map_struct(maps)), float(bias))


def clean_mask(canny):
    canny, p = _u8(canny)
    h, w = canny.shape
    out = np.empty((h, w), np.uint8)
    lib().mo_clean_mask(p, w, h, _dp(out))
    return out


def render_grid(R_w2c, t_w2c, K, D, w, h, line_width=0.04, floor_level=30):
    R = np.ascontiguousarray(R_w2c, np.float64)
    t = np.ascontiguousarray(t_w2c, np.float64)
    K = np.ascontiguousarray(K, np.float32)
    D = np.ascontiguousarray(D, np.float64)
    out = np.empty((h, w, 3), np.uint8)
    f = lib().mo_render_grid
    f.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_int, ctypes.c_int, ctypes.c_double, ctypes.c_int, ctypes.c_void_p]
    f(_dp(R), _dp(t), _dp(K), _dp(D), w, h, float(line_width), int(floor_level), _dp(out))
    return out


# ---- Hough (row a6'; cv::HoughLines semantics, see oracle/hough.c) ----
class HoughLine(ctypes.Structure):
    _fields_ = [("rho", ctypes.c_float), ("theta", ctypes.c_float), ("votes", ctypes.c_int)]


def hough_dims(w, h, rho=1.0, theta=np.pi / 180):
    na, nr = ctypes.c_int(), ctypes.c_int()
    f = lib().mo_hough_dims
    f.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_double, ctypes.c_double, ctypes.c_void_p, ctypes.c_void_p]
    f(w, h, float(rho), float(theta), ctypes.byref(na), ctypes.byref(nr))
    return na.value, nr.value


def hough_lines(edges, threshold, rho=1.0, theta=np.pi / 180, max_lines=4096, want_accum=False):
    """Returns (lines [n,3] float64 of rho, theta, votes) and optionally the accumulator."""
    edges, p = _u8(edges)
    h, w = edges.shape
    out = (HoughLine * max_lines)()
    na, nr = hough_dims(w, h, rho, theta)
    acc = np.zeros((na + 2, nr + 2), np.int32) if want_accum else None
    f = lib().mo_hough_lines
    f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_double, ctypes.c_double, ctypes.c_int,
                  ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    n = f(p, w, h, float(rho), float(theta), int(threshold), ctypes.addressof(out), max_lines,
          acc.ctypes.data if want_accum else None)
    lines = np.array([[out[i].rho, out[i].theta, out[i].votes] for i in range(n)], np.float64).reshape(n, 3)
    return (lines, acc) if want_accum else lines


class HoughCluster(ctypes.Structure):
    _fields_ = [("rho", ctypes.c_float), ("theta", ctypes.c_float), ("votes", ctypes.c_int), ("members", ctypes.c_int)]


def hough_clusters(lines, w, h, rho_tol=8.0, theta_tol=np.deg2rad(4.0), theta=np.pi / 180, max_clusters=256):
    """lines: [n,3] (rho, theta, votes) strongest first -> [m,4] (rho, theta, votes, members)."""
    n = len(lines)
    arr = (HoughLine * max(n, 1))()
    for i in range(n):
        arr[i].rho, arr[i].theta, arr[i].votes = float(lines[i][0]), float(lines[i][1]), int(lines[i][2])
    out = (HoughCluster * max_clusters)()
    na, _ = hough_dims(w, h, 1.0, theta)
    f = lib().mo_hough_cluster_lines
    f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_double, ctypes.c_int, ctypes.c_double, ctypes.c_double,
                  ctypes.c_void_p, ctypes.c_int]
    m = f(ctypes.addressof(arr), n, float(theta), na, float(rho_tol), float(theta_tol), ctypes.addressof(out), max_clusters)
    return np.array([[out[i].rho, out[i].theta, out[i].votes, out[i].members] for i in range(m)], np.float64).reshape(m, 4)


def hough_intersections(clusters, w, h, pencil_tol=np.deg2rad(25.0), max_points=1024):
    n = len(clusters)
    arr = (HoughCluster * max(n, 1))()
    for i in range(n):
        arr[i].rho, arr[i].theta = float(clusters[i][0]), float(clusters[i][1])
        arr[i].votes, arr[i].members = int(clusters[i][2]), int(clusters[i][3])
    xy = np.zeros((max_points, 2), np.float32)
    pen = np.zeros(max(n, 1), np.int32)
    f = lib().mo_hough_intersections
    f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_double, ctypes.c_void_p,
                  ctypes.c_int, ctypes.c_void_p]
    m = f(ctypes.addressof(arr), n, w, h, float(pencil_tol), xy.ctypes.data, max_points, pen.ctypes.data)
    return xy[:m].copy(), pen[:n].copy()


# ---- PnP (config 5; cv::solvePnP(ITERATIVE) semantics, see oracle/pnp.c) ----
def pnp_iterative(obj, img, K, guess=None, both_planar=False, max_iters=50):
    """obj [m,3], img [m,2] (converted to float32 as the C ABI takes them) -> dict(rvec, tvec, err, passes, ok)."""
    obj = np.ascontiguousarray(obj, np.float32)
    img = np.ascontiguousarray(img, np.float32)
    Kf = np.ascontiguousarray(K, np.float32).reshape(-1)
    g = None if guess is None else np.ascontiguousarray(guess, np.float64).reshape(6)
    out = np.zeros(9, np.float64)
    f = lib().mo_pnp_iterative
    f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                  ctypes.c_int, ctypes.c_void_p]
    ok = f(obj.ctypes.data, img.ctypes.data, len(obj), Kf.ctypes.data, None if g is None else g.ctypes.data,
           1 if both_planar else 0, int(max_iters), out.ctypes.data)
    return dict(rvec=out[:3].copy(), tvec=out[3:6].copy(), err=float(out[6]), passes=int(out[7]), ok=bool(ok))


def rodrigues_exp(w):
    R = np.zeros(9)
    f = lib().mo_rodrigues_exp
    f.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    w = np.ascontiguousarray(w, np.float64)
    f(w.ctypes.data, R.ctypes.data)
    return R.reshape(3, 3)


# ---- SURVEY 8(f) rows 3-4 (oracle/legacy.c) ----
def undistort_image(bgr, K, D):
    """cv::fisheye::undistortImage(img, K, D, Knew=K) (MonteCarlo.cpp:284)."""
    bgr, p = _u8(bgr)
    h, w, _ = bgr.shape
    Kf = np.ascontiguousarray(K, np.float32).reshape(-1)
    Dd = np.ascontiguousarray(D, np.float64).reshape(-1)
    m1 = np.zeros((h, w, 2), np.int16)
    m2 = np.zeros((h, w), np.uint16)
    f = lib().mo_fisheye_undistort_maps
    f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    f(Kf.ctypes.data, Dd.ctypes.data, w, h, m1.ctypes.data, m2.ctypes.data)
    out = np.zeros_like(bgr)
    g = lib().mo_remap_linear_bgr
    g.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                  ctypes.c_int]
    g(p, w, h, 3 * w, m1.ctypes.data, m2.ctypes.data, out.ctypes.data, 3 * w)
    return out, m1, m2


def camera_error(bgr, K, w2c12, white, red, green):
    """MonteCarlo::computeCameraError (MonteCarlo.cpp:183-226); w2c12 = R row-major (9) + t (3)."""
    bgr, p = _u8(bgr)
    h, w, _ = bgr.shape
    Kf = np.ascontiguousarray(K, np.float32).reshape(-1)
    T = np.ascontiguousarray(w2c12, np.float64).reshape(12)
    mp = np.ascontiguousarray(np.concatenate([white, red, green]), np.float64)
    f = lib().mo_camera_error
    f.restype = ctypes.c_double
    f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                  ctypes.c_int, ctypes.c_int, ctypes.c_int]
    return f(p, w, h, 3 * w, Kf.ctypes.data, T.ctypes.data, mp.ctypes.data, len(white), len(red), len(green))


class Markov:
    """MarkovModel (include/mantis3/Markov.cpp) on one 360-bin plane."""

    def __init__(self, yaw_rad=None, plane=None):
        self.p = np.zeros(360, np.float64) if plane is None else np.ascontiguousarray(plane, np.float64).copy()
        if yaw_rad is not None:
            self._call("mo_markov_init", float(yaw_rad))

    def _call(self, name, *a):
        f = getattr(lib(), name)
        f.argtypes = [ctypes.c_void_p] + [ctypes.c_double] * len(a)
        f.restype = None
        f(self.p.ctypes.data, *a)

    def update_weights(self, stddev):
        self._call("mo_markov_update_weights", float(stddev))

    def convolve(self, dtheta, dt):
        self._call("mo_markov_convolve", float(dtheta), float(dt))

    def sense(self, yaw_rad):
        self._call("mo_markov_sense", float(yaw_rad))

    def update_errors(self, yaw_rad, err):
        y = np.ascontiguousarray(yaw_rad, np.float64)
        e = np.ascontiguousarray(err, np.float64).copy()
        f = lib().mo_markov_update_errors
        f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
        f.restype = None
        f(self.p.ctypes.data, y.ctypes.data, e.ctypes.data, len(e))
        return e

    def get_yaw(self):
        b = ctypes.c_int()
        f = lib().mo_markov_get_yaw
        f.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        f.restype = ctypes.c_double
        return f(self.p.ctypes.data, ctypes.byref(b)), b.value
