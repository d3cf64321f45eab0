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
