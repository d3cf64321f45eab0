"""ctypes binding of libmantis_cuda.so (include/mantis_cuda.h).

Mirrors the reference's stage functions by name where one exists
(``detectQuadrilaterals`` image stage -> :meth:`Context.edges`, ...).  Fails loudly if the
shared library is missing; there is no Python / CPU implementation behind it.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

MEM_HOST = 0
MEM_DEVICE = 1


class MantisError(RuntimeError):
    pass


class Config(ctypes.Structure):
    """mantis_config: rosparams + defines of include/mantis3/Mantis3Params.h."""

    _fields_ = [
        ("canny_hysteresis", ctypes.c_int32),
        ("polygon_epsilon", ctypes.c_int32),
        ("search_radius_multiplier", ctypes.c_double),
        ("grid_spacing", ctypes.c_double),
        ("grid_size", ctypes.c_int32),
        ("quad_stretch_factor", ctypes.c_double),
        ("max_quad_error", ctypes.c_double),
        ("max_angle_difference", ctypes.c_double),
        ("projection_bias", ctypes.c_double),
        ("color_projection_bias", ctypes.c_double),
        ("rot_sigma", ctypes.c_double),
        ("trans_sigma", ctypes.c_double),
        ("particles", ctypes.c_int32),
        ("iterations", ctypes.c_int32),
        ("rng_seed", ctypes.c_uint64),
    ]


MAX_QUADS = 256


class FrameQuads(ctypes.Structure):
    """mantis_frame_quads."""

    _fields_ = [("n_contours", ctypes.c_int32), ("n_quads_raw", ctypes.c_int32), ("n_quads", ctypes.c_int32),
                ("flags", ctypes.c_int32), ("corners", (ctypes.c_int32 * 8) * MAX_QUADS),
                ("test_px", (ctypes.c_double * 8) * MAX_QUADS)]


class Frame(ctypes.Structure):
    """mantis_frame: one Image + CameraInfo of mantisService.srv."""

    _fields_ = [("bgr", ctypes.c_void_p), ("width", ctypes.c_int32), ("height", ctypes.c_int32),
                ("stride_bytes", ctypes.c_int32), ("K", ctypes.c_float * 9), ("D", ctypes.c_double * 4),
                ("cam_to_base", ctypes.c_double * 7), ("stamp_ns", ctypes.c_uint64), ("camera_id", ctypes.c_int32)]


class Result(ctypes.Structure):
    """mantis_result: pose + error/weight + counts (PosePub.h:38-56, mantisService.srv:9-12)."""

    _fields_ = [("pos", ctypes.c_double * 3), ("quat_xyzw", ctypes.c_double * 4), ("error", ctypes.c_double),
                ("min_yaw_diff", ctypes.c_double), ("status", ctypes.c_int32), ("n_quads", ctypes.c_int32),
                ("n_hyps", ctypes.c_int32), ("n_particles", ctypes.c_int32)]


class FrameTrace(ctypes.Structure):
    """mantis_frame_trace."""

    _fields_ = [("n_contours", ctypes.c_int32), ("n_quads_raw", ctypes.c_int32), ("n_quads", ctypes.c_int32),
                ("flags", ctypes.c_int32), ("n_hyps_generated", ctypes.c_int32), ("n_hyps_cluster", ctypes.c_int32),
                ("best_rot", ctypes.c_int32), ("pad_", ctypes.c_int32), ("first_best_error", ctypes.c_double),
                ("optimized_error", ctypes.c_double), ("optimized_w2c", ctypes.c_double * 12),
                ("shift_err", ctypes.c_double * 81), ("yaw_err", ctypes.c_double * 4)]


HOUGH_MAX_LINES, HOUGH_MAX_CLUSTERS, HOUGH_MAX_POINTS = 1024, 256, 1024


class HoughLine(ctypes.Structure):
    _fields_ = [("rho", ctypes.c_float), ("theta", ctypes.c_float), ("votes", ctypes.c_int32)]


class HoughCluster(ctypes.Structure):
    _fields_ = [("rho", ctypes.c_float), ("theta", ctypes.c_float), ("votes", ctypes.c_int32),
                ("members", ctypes.c_int32)]


class FrameLines(ctypes.Structure):
    """mantis_frame_lines."""

    _fields_ = [("n_peaks", ctypes.c_int32), ("n_lines", ctypes.c_int32), ("n_clusters", ctypes.c_int32),
                ("n_points", ctypes.c_int32), ("flags", ctypes.c_int32), ("pad_", ctypes.c_int32),
                ("lines", HoughLine * HOUGH_MAX_LINES), ("clusters", HoughCluster * HOUGH_MAX_CLUSTERS),
                ("points", (ctypes.c_float * 2) * HOUGH_MAX_POINTS)]


class HoughParams(ctypes.Structure):
    """mantis_hough_params."""

    _fields_ = [("threshold", ctypes.c_int32), ("use_morph", ctypes.c_int32), ("rho_tol", ctypes.c_double),
                ("theta_tol", ctypes.c_double), ("pencil_tol", ctypes.c_double)]


class PnpParams(ctypes.Structure):
    _fields_ = [("flags", ctypes.c_int32), ("max_iters", ctypes.c_int32)]


class PnpResult(ctypes.Structure):
    """mantis_pnp_result."""

    _fields_ = [("rvec", ctypes.c_double * 3), ("tvec", ctypes.c_double * 3), ("reproj_error", ctypes.c_double),
                ("passes", ctypes.c_int32), ("status", ctypes.c_int32)]


class RppResult(ctypes.Structure):
    """mantis_rpp_result."""

    _fields_ = [("R", ctypes.c_double * 9), ("t", ctypes.c_double * 3), ("obj_err", ctypes.c_double),
                ("img_err", ctypes.c_double), ("iters_first", ctypes.c_int32), ("iters_best", ctypes.c_int32),
                ("n_candidates", ctypes.c_int32), ("status", ctypes.c_int32)]


PNP_BOTH_PLANAR = 1
NOALLOC_SCRATCH_BYTES = 8 << 20  # MANTIS_NOALLOC_SCRATCH_BYTES


def library_path():
    return os.path.join(_HERE, "libmantis_cuda.so")


def build_library(force=False):
    """Compile csrc/*.cu for sm_100a into mantis_b200/libmantis_cuda.so (in-tree)."""
    args = ["make", "-C", os.path.join(_HERE, "csrc"), "-j8", "-s"]
    if force:
        args.append("-B")
    subprocess.check_call(args)
    return library_path()


def load_library():
    global _LIB
    if _LIB is not None:
        return _LIB
    path = library_path()
    if not os.path.exists(path):
        raise MantisError(
            f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(mantis_b200 has no CPU fallback)"
        )
    lib = ctypes.CDLL(path)
    c = ctypes
    lib.mantis_default_config.argtypes = [c.POINTER(Config)]
    lib.mantis_default_config.restype = None
    lib.mantis_create.argtypes = [c.POINTER(Config), c.c_int, c.POINTER(c.c_void_p)]
    lib.mantis_destroy.argtypes = [c.c_void_p]
    lib.mantis_strerror.argtypes = [c.c_int]
    lib.mantis_strerror.restype = c.c_char_p
    lib.mantis_last_error.argtypes = [c.c_void_p]
    lib.mantis_last_error.restype = c.c_char_p
    lib.mantis_launch_count.argtypes = [c.c_void_p]
    lib.mantis_launch_count.restype = c.c_uint64
    lib.mantis_stream.argtypes = [c.c_void_p]
    lib.mantis_stream.restype = c.c_void_p
    lib.mantis_edges_batch.argtypes = [c.c_void_p, c.c_void_p, c.c_int, c.c_int, c.c_int, c.c_int, c.c_size_t,
                                       c.c_void_p, c.c_void_p, c.c_int]
    lib.mantis_mask_batch.argtypes = [c.c_void_p, c.c_void_p, c.c_int, c.c_int, c.c_int, c.c_int, c.c_size_t,
                                      c.c_void_p, c.c_int]
    lib.mantis_quads_batch.argtypes = [c.c_void_p, c.c_void_p, c.c_int, c.c_int, c.c_int, c.c_int, c.c_size_t,
                                       c.c_void_p, c.c_int]
    lib.mantis_hough_batch.argtypes = [c.c_void_p, c.c_void_p, c.c_int, c.c_int, c.c_int, c.c_int, c.c_size_t,
                                       c.c_void_p, c.c_void_p, c.c_int]
    lib.mantis_pnp_batch.argtypes = [c.c_void_p, c.c_void_p, c.c_void_p, c.c_int, c.c_int, c.c_int, c.c_void_p, c.c_int,
                                     c.c_void_p, c.c_void_p, c.c_void_p, c.c_int]
    lib.mantis_particle_weights_batch.argtypes = [c.c_void_p, c.c_void_p, c.c_int, c.c_int, c.c_int, c.c_int, c.c_size_t,
                                                  c.c_void_p, c.c_void_p, c.c_int, c.c_void_p, c.c_int]
    lib.mantis_undistort_batch.argtypes = [c.c_void_p, c.c_void_p, c.c_int, c.c_int, c.c_int, c.c_int, c.c_size_t,
                                           c.c_void_p, c.c_void_p, c.c_void_p, c.c_int]
    lib.mantis_markov_batch.argtypes = [c.c_void_p, c.c_void_p, c.c_int, c.c_int, c.c_void_p, c.c_void_p, c.c_int]
    lib.mantis_markov_update_errors.argtypes = [c.c_void_p, c.c_void_p, c.c_int, c.c_void_p, c.c_void_p, c.c_void_p,
                                                c.c_int, c.c_int]
    lib.mantis_set_map.argtypes = [c.c_void_p, c.c_void_p, c.c_int, c.c_int, c.c_int]
    lib.mantis_localize_batch.argtypes = [c.c_void_p, c.c_void_p, c.c_int, c.c_void_p, c.c_void_p, c.c_int]
    lib.mantis_localize_batch_async.argtypes = [c.c_void_p, c.c_void_p, c.c_int, c.c_void_p, c.c_void_p, c.c_int]
    lib.mantis_wait.argtypes = [c.c_void_p]
    lib.mantis_localize_rig_batch.argtypes = [c.c_void_p, c.c_void_p, c.c_int, c.c_int, c.c_void_p, c.c_void_p, c.c_int]
    lib.mantis_render_grid_batch.argtypes = [c.c_void_p, c.c_void_p, c.c_int, c.c_int, c.c_int, c.c_void_p,
                                             c.c_void_p, c.c_double, c.c_int, c.c_float, c.c_uint32, c.c_void_p,
                                             c.c_int, c.c_size_t]
    lib.mantis_degrade_batch.argtypes = [c.c_void_p, c.c_void_p, c.c_void_p, c.c_int, c.c_int, c.c_int, c.c_int, c.c_size_t,
                                         c.c_int, c.c_size_t, c.c_float, c.c_float, c.c_uint32]
    lib.mantis_rpp_batch.argtypes = [c.c_void_p, c.c_void_p, c.c_int, c.c_void_p, c.c_int, c.c_int, c.c_void_p, c.c_int]
    lib.mantis_score_batch.argtypes = [c.c_void_p, c.c_void_p, c.c_int, c.c_int, c.c_int, c.c_int, c.c_size_t, c.c_void_p,
                                       c.c_void_p, c.c_void_p, c.c_int, c.c_void_p, c.c_int]
    lib.mantis_edges_noalloc.argtypes = [c.c_void_p, c.c_void_p, c.c_int, c.c_int, c.c_int, c.c_int, c.c_size_t, c.c_void_p,
                                         c.c_void_p]
    lib.mantis_set_profiling.argtypes = [c.c_void_p, c.c_int]
    lib.mantis_get_stage_ms.argtypes = [c.c_void_p, c.c_void_p]
    _LIB = lib
    return lib


def default_config():
    cfg = Config()
    load_library().mantis_default_config(ctypes.byref(cfg))
    return cfg


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, int):
        return ctypes.c_void_p(a)
    return ctypes.c_void_p(a.ctypes.data)


class Context:
    """One mantis_ctx: bound to a device and a stream; not thread-safe."""

    def __init__(self, device=0, config=None):
        self._lib = load_library()
        self.config = config if config is not None else default_config()
        h = ctypes.c_void_p()
        rc = self._lib.mantis_create(ctypes.byref(self.config), int(device), ctypes.byref(h))
        if rc != 0:
            raise MantisError(f"mantis_create: {self._lib.mantis_strerror(rc).decode()}")
        self._h = h
        self.device = device

    def close(self):
        if getattr(self, "_h", None):
            self._lib.mantis_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc, what):
        if rc != 0:
            raise MantisError(f"{what}: {self._lib.mantis_strerror(rc).decode()} "
                              f"[{self._lib.mantis_last_error(self._h).decode()}]")

    @property
    def launch_count(self):
        return int(self._lib.mantis_launch_count(self._h))

    @property
    def stream(self):
        return int(self._lib.mantis_stream(self._h) or 0)

    # -- image stage: detectQuadrilaterals(), QuadDetection.h:209-214 --
    def edges(self, bgr, want_canny=True, want_morph=True):
        """bgr: uint8 array [n, H, W, 3] (or [H, W, 3]) on the host.  Returns (canny, morph)."""
        bgr = np.ascontiguousarray(bgr, dtype=np.uint8)
        if bgr.ndim == 3:
            bgr = bgr[None]
        n, h, w, ch = bgr.shape
        assert ch == 3
        canny = np.empty((n, h, w), np.uint8) if want_canny else None
        morph = np.empty((n, h, w), np.uint8) if want_morph else None
        rc = self._lib.mantis_edges_batch(self._h, _ptr(bgr), n, w, h, 3 * w, 3 * w * h, _ptr(canny), _ptr(morph),
                                          MEM_HOST)
        self._check(rc, "mantis_edges_batch")
        return canny, morph

    def edges_device(self, d_bgr, n, w, h, d_canny=None, d_morph=None, stride=None, pitch=None):
        """Device-resident variant: integer device pointers (e.g. torch .data_ptr())."""
        stride = 3 * w if stride is None else stride
        pitch = stride * h if pitch is None else pitch
        rc = self._lib.mantis_edges_batch(self._h, _ptr(d_bgr), n, w, h, stride, pitch, _ptr(d_canny), _ptr(d_morph),
                                          MEM_DEVICE)
        self._check(rc, "mantis_edges_batch")

    def edges_noalloc(self, d_bgr, n, w, h, d_canny=None, d_morph=None, stride=None, pitch=None):
        """mantis_edges_noalloc: the device-resident image stage with its scratch in a fixed __device__ array
        (no cudaMalloc inside the call; at most NOALLOC_SCRATCH_BYTES of bit planes per call)."""
        stride = 3 * w if stride is None else stride
        pitch = stride * h if pitch is None else pitch
        rc = self._lib.mantis_edges_noalloc(self._h, _ptr(d_bgr), n, w, h, stride, pitch, _ptr(d_canny), _ptr(d_morph))
        self._check(rc, "mantis_edges_noalloc")

    # -- cleanImageByEdge mask, HypothesisEvaluation.h:319-386 --
    def edge_mask(self, bgr):
        bgr = np.ascontiguousarray(bgr, dtype=np.uint8)
        if bgr.ndim == 3:
            bgr = bgr[None]
        n, h, w, _ = bgr.shape
        mask = np.empty((n, h, w), np.uint8)
        rc = self._lib.mantis_mask_batch(self._h, _ptr(bgr), n, w, h, 3 * w, 3 * w * h, _ptr(mask), MEM_HOST)
        self._check(rc, "mantis_mask_batch")
        return mask

    # -- detectQuadrilaterals(), QuadDetection.h:203-287 --
    def quads(self, bgr):
        """Returns one FrameQuads record per frame."""
        bgr = np.ascontiguousarray(bgr, dtype=np.uint8)
        if bgr.ndim == 3:
            bgr = bgr[None]
        n, h, w, _ = bgr.shape
        out = (FrameQuads * n)()
        rc = self._lib.mantis_quads_batch(self._h, _ptr(bgr), n, w, h, 3 * w, 3 * w * h, ctypes.addressof(out), MEM_HOST)
        self._check(rc, "mantis_quads_batch")
        return out

    # -- line voting (north-star stage; cv::HoughLines semantics, no reference call site) --
    def hough(self, bgr, threshold, use_morph=False, rho_tol=8.0, theta_tol=np.deg2rad(4.0),
              pencil_tol=np.deg2rad(25.0)):
        """Returns one FrameLines record per frame (peaks, clustered lines, pencil crossings)."""
        bgr = np.ascontiguousarray(bgr, dtype=np.uint8)
        if bgr.ndim == 3:
            bgr = bgr[None]
        n, h, w, _ = bgr.shape
        out = (FrameLines * n)()
        prm = HoughParams(int(threshold), int(use_morph), float(rho_tol), float(theta_tol), float(pencil_tol))
        rc = self._lib.mantis_hough_batch(self._h, _ptr(bgr), n, w, h, 3 * w, 3 * w * h, ctypes.byref(prm),
                                          ctypes.addressof(out), MEM_HOST)
        self._check(rc, "mantis_hough_batch")
        return out

    # -- computeHypothesis(): solvePnP(ITERATIVE) + reprojection error, legacy/mantis2/PoseEstimator.h:16-25,91-118 --
    def pnp(self, obj, img, K, guess=None, both_planar=False, max_iters=0):
        """obj [n,m,3] or [m,3] (shared), img [n,m,2], K [3,3] (shared) or [n,3,3] -> (PnpResult * n)."""
        img = np.ascontiguousarray(img, np.float32)
        obj = np.ascontiguousarray(obj, np.float32)
        K = np.ascontiguousarray(K, np.float32)
        n, m, _ = img.shape
        g = None if guess is None else np.ascontiguousarray(guess, np.float64).reshape(n, 6)
        out = (PnpResult * n)()
        prm = PnpParams(PNP_BOTH_PLANAR if both_planar else 0, int(max_iters))
        rc = self._lib.mantis_pnp_batch(self._h, _ptr(obj), _ptr(img), n, m, int(obj.ndim == 2), _ptr(K), int(K.ndim == 2),
                                        _ptr(g), ctypes.byref(prm), ctypes.addressof(out), MEM_HOST)
        self._check(rc, "mantis_pnp_batch")
        return out

    def pnp_device(self, d_obj, d_img, n, m, d_K, d_out, obj_shared=False, k_shared=True, d_guess=None, both_planar=False,
                   max_iters=0):
        """Device-resident variant (integer device pointers); d_out receives n mantis_pnp_result records."""
        prm = PnpParams(PNP_BOTH_PLANAR if both_planar else 0, int(max_iters))
        rc = self._lib.mantis_pnp_batch(self._h, _ptr(d_obj), _ptr(d_img), n, m, int(obj_shared), _ptr(d_K), int(k_shared),
                                        _ptr(d_guess), ctypes.byref(prm), _ptr(d_out), MEM_DEVICE)
        self._check(rc, "mantis_pnp_batch")

    # -- RPP::Rpp per problem, RobustPlanarPose/RPP.h:54-55 (CoPlanarPoseEstimator.cpp:16-58) --
    def rpp(self, model_xyz, image_xy):
        """model_xyz [n,m,3] or [m,3] (shared), image_xy [n,m,2] normalised image points -> (RppResult * n)."""
        model = np.ascontiguousarray(model_xyz, np.float64)
        img = np.ascontiguousarray(image_xy, np.float64)
        n, m, _ = img.shape
        out = (RppResult * n)()
        rc = self._lib.mantis_rpp_batch(self._h, _ptr(model), int(model.ndim == 2), _ptr(img), n, m, ctypes.addressof(out),
                                        MEM_HOST)
        self._check(rc, "mantis_rpp_batch")
        return out

    # -- evaluateHypothesisWithImageWHITE on cleanImageByEdge's image, HypothesisEvaluation.h:107-158, 319-386 --
    def score(self, bgr, K, D, w2c):
        """bgr [n,H,W,3], K [3,3] or [n,3,3], D [4] or [n,4], w2c [n,P,12] (R row-major + t) -> scores [n,P]."""
        bgr = np.ascontiguousarray(bgr, np.uint8)
        if bgr.ndim == 3:
            bgr = bgr[None]
        n, h, w, _ = bgr.shape
        K = np.ascontiguousarray(K, np.float32)
        if K.ndim == 2:
            K = np.repeat(K[None], n, 0)
        D = np.ascontiguousarray(D, np.float64)
        if D.ndim == 1:
            D = np.repeat(D[None], n, 0)
        w2c = np.ascontiguousarray(w2c, np.float64).reshape(n, -1, 12)
        out = np.zeros((n, w2c.shape[1]), np.float64)
        rc = self._lib.mantis_score_batch(self._h, _ptr(bgr), n, w, h, 3 * w, 3 * w * h, _ptr(K), _ptr(D), _ptr(w2c),
                                          w2c.shape[1], _ptr(out), MEM_HOST)
        self._check(rc, "mantis_score_batch")
        return out

    # -- 2016 generation: computeCameraError / undistortImage, legacy/mantis/MonteCarlo.cpp:169-241,284 --
    def particle_weights(self, bgr, K, w2c):
        """bgr [n,H,W,3] (undistorted), K [n,3,3] or [3,3], w2c [n,P,12] -> weights [n,P]."""
        bgr = np.ascontiguousarray(bgr, np.uint8)
        if bgr.ndim == 3:
            bgr = bgr[None]
        n, h, w, _ = bgr.shape
        K = np.ascontiguousarray(K, np.float32)
        if K.ndim == 2:
            K = np.repeat(K[None], n, 0)
        w2c = np.ascontiguousarray(w2c, np.float64).reshape(n, -1, 12)
        out = np.zeros((n, w2c.shape[1]), np.float64)
        rc = self._lib.mantis_particle_weights_batch(self._h, _ptr(bgr), n, w, h, 3 * w, 3 * w * h, _ptr(K), _ptr(w2c),
                                                     w2c.shape[1], _ptr(out), MEM_HOST)
        self._check(rc, "mantis_particle_weights_batch")
        return out

    def undistort(self, bgr, K, D):
        bgr = np.ascontiguousarray(bgr, np.uint8)
        if bgr.ndim == 3:
            bgr = bgr[None]
        n, h, w, _ = bgr.shape
        Kf, Dd = np.ascontiguousarray(K, np.float32), np.ascontiguousarray(D, np.float64)
        out = np.zeros_like(bgr)
        rc = self._lib.mantis_undistort_batch(self._h, _ptr(bgr), n, w, h, 3 * w, 3 * w * h, _ptr(Kf), _ptr(Dd), _ptr(out),
                                              MEM_HOST)
        self._check(rc, "mantis_undistort_batch")
        return out

    # -- MarkovModel, include/mantis3/Markov.cpp --
    MARKOV_INIT, MARKOV_CONVOLVE, MARKOV_SENSE = 0, 1, 2

    def markov(self, planes, op, a, b=None):
        """planes [S,360] float64 (updated copy returned); a, b per-stream arguments of the operation."""
        planes = np.ascontiguousarray(planes, np.float64).copy()
        a = np.ascontiguousarray(a, np.float64)
        b = None if b is None else np.ascontiguousarray(b, np.float64)
        rc = self._lib.mantis_markov_batch(self._h, _ptr(planes), len(planes), int(op), _ptr(a), _ptr(b), MEM_HOST)
        self._check(rc, "mantis_markov_batch")
        return planes

    def markov_update_errors(self, planes, stream_of, yaw, err):
        planes = np.ascontiguousarray(planes, np.float64)
        so = np.ascontiguousarray(stream_of, np.int32)
        yaw = np.ascontiguousarray(yaw, np.float64)
        err = np.ascontiguousarray(err, np.float64).copy()
        rc = self._lib.mantis_markov_update_errors(self._h, _ptr(planes), len(planes), _ptr(so), _ptr(yaw), _ptr(err),
                                                   len(err), MEM_HOST)
        self._check(rc, "mantis_markov_update_errors")
        return err

    # -- quadDetection(), src/mantis3.cpp:68-135 --
    def set_map(self, white, red, green):
        xyz = np.ascontiguousarray(np.concatenate([white, red, green]), np.float64)
        rc = self._lib.mantis_set_map(self._h, _ptr(xyz), len(white), len(red), len(green))
        self._check(rc, "mantis_set_map")

    def make_frames(self, ptrs, w, h, K, D, stride=None):
        """Array of mantis_frame for buffers (host arrays or device pointers) sharing K / D."""
        n = len(ptrs)
        fr = (Frame * n)()
        Kf = np.ascontiguousarray(K, np.float32).reshape(-1)
        Dd = np.ascontiguousarray(D, np.float64).reshape(-1)
        for i, p in enumerate(ptrs):
            fr[i].bgr = p if isinstance(p, int) else p.ctypes.data
            fr[i].width, fr[i].height, fr[i].stride_bytes = w, h, (3 * w if stride is None else stride)
            fr[i].K[:] = Kf.tolist()
            fr[i].D[:] = Dd.tolist()
            fr[i].camera_id = 0
        return fr

    def localize(self, bgr, K, D, want_trace=False):
        """bgr: uint8 [n, H, W, 3] on the host, one camera model for all.  Returns (results, traces)."""
        bgr = np.ascontiguousarray(bgr, dtype=np.uint8)
        if bgr.ndim == 3:
            bgr = bgr[None]
        n, h, w, _ = bgr.shape
        frames = self.make_frames([bgr[i] for i in range(n)], w, h, K, D)
        return self.localize_frames(frames, MEM_HOST, want_trace)

    # -- rigs: joint multi-camera localization (MantisRequest.h:14-26, Mantis2Error.h:55-99) --
    def localize_rig(self, bgr, Ks, Ds, cam_to_base, want_trace=False):
        """bgr [R, C, H, W, 3] (R rigs of C cameras) on the host; Ks [C,3,3], Ds [C,4], cam_to_base [C,7]
        (x y z qx qy qz qw of each camera in the base frame) shared by all rigs.  One Result per rig (base pose)."""
        bgr = np.ascontiguousarray(bgr, dtype=np.uint8)
        if bgr.ndim == 4:
            bgr = bgr[None]
        R, C, h, w, _ = bgr.shape
        fr = (Frame * (R * C))()
        Ks = np.ascontiguousarray(Ks, np.float32).reshape(C, 9)
        Ds = np.ascontiguousarray(Ds, np.float64).reshape(C, 4)
        b2c = np.ascontiguousarray(cam_to_base, np.float64).reshape(C, 7)
        for r in range(R):
            for k in range(C):
                f = fr[r * C + k]
                f.bgr = bgr[r, k].ctypes.data
                f.width, f.height, f.stride_bytes = w, h, 3 * w
                f.K[:] = Ks[k].tolist()
                f.D[:] = Ds[k].tolist()
                f.cam_to_base[:] = b2c[k].tolist()
                f.camera_id = k
        return self.localize_rig_frames(fr, R, C, MEM_HOST, want_trace)

    def localize_rig_frames(self, frames, n_rigs, cams_per_rig, mem, want_trace=False):
        out = (Result * n_rigs)()
        tr = (FrameTrace * n_rigs)() if want_trace else None
        rc = self._lib.mantis_localize_rig_batch(self._h, ctypes.addressof(frames), n_rigs, cams_per_rig, ctypes.addressof(out),
                                                 ctypes.addressof(tr) if want_trace else None, mem)
        self._check(rc, "mantis_localize_rig_batch")
        return out, tr

    def localize_frames(self, frames, mem, want_trace=False):
        n = len(frames)
        out = (Result * n)()
        tr = (FrameTrace * n)() if want_trace else None
        rc = self._lib.mantis_localize_batch(self._h, ctypes.addressof(frames), n, ctypes.addressof(out),
                                             ctypes.addressof(tr) if want_trace else None, mem)
        self._check(rc, "mantis_localize_batch")
        return out, tr

    def localize_frames_async(self, frames, mem, out, trace=None):
        """mantis_localize_batch_async: queue the batch; `out` (Result array, ideally in pinned memory) is filled once
        wait() has returned.  Keep `frames` and `out` alive until then."""
        rc = self._lib.mantis_localize_batch_async(self._h, ctypes.addressof(frames), len(frames), ctypes.addressof(out),
                                                   ctypes.addressof(trace) if trace is not None else None, mem)
        self._check(rc, "mantis_localize_batch_async")

    def wait(self):
        self._check(self._lib.mantis_wait(self._h), "mantis_wait")

    # -- synthetic frame source (stands in for GridRenderer.cpp:292-353) --
    def render_grid_device(self, poses, w, h, K, D, d_out, line_width=0.04, floor_level=30, noise_sigma=0.0, seed=0):
        """poses: list of (R_w2c[3,3], t_w2c[3]); d_out: device pointer to n*h*w*3 bytes."""
        flat = np.ascontiguousarray([np.concatenate([np.asarray(R, np.float64).reshape(-1), np.asarray(t, np.float64)])
                                     for R, t in poses], np.float64)
        Kf = np.ascontiguousarray(K, np.float32)
        Dd = np.ascontiguousarray(D, np.float64)
        rc = self._lib.mantis_render_grid_batch(self._h, _ptr(flat), len(poses), w, h, _ptr(Kf), _ptr(Dd),
                                                float(line_width), int(floor_level), float(noise_sigma), int(seed),
                                                _ptr(d_out), 3 * w, 3 * w * h)
        self._check(rc, "mantis_render_grid_batch")

    def degrade_device(self, d_in, d_out, n, w, h, blur_sigma=1.5, noise_sigma=8.0, seed=0):
        """BASELINE configs[3]'s blur + noise on device-resident frames (tightly packed); d_in != d_out."""
        rc = self._lib.mantis_degrade_batch(self._h, _ptr(d_in), _ptr(d_out), n, w, h, 3 * w, 3 * w * h, 3 * w, 3 * w * h,
                                            float(blur_sigma), float(noise_sigma), int(seed))
        self._check(rc, "mantis_degrade_batch")

    STAGES = ("h2d", "edge_candidates", "edge_link", "quads", "pose", "search", "d2h")

    def set_profiling(self, enabled=True):
        self._check(self._lib.mantis_set_profiling(self._h, int(enabled)), "mantis_set_profiling")

    def stage_ms(self):
        """Accumulated device milliseconds per stage since the last call (dict)."""
        ms = (ctypes.c_double * 7)()
        self._check(self._lib.mantis_get_stage_ms(self._h, ms), "mantis_get_stage_ms")
        return dict(zip(self.STAGES, list(ms)))
