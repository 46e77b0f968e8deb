"""ctypes loader for oracle/_build/libabub_ref.so (TEST INFRASTRUCTURE, see oracle/__init__.py)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "libabub_ref.so")


class RefBlob(C.Structure):
    _fields_ = [("k", C.c_int32), ("x", C.c_int32), ("y", C.c_int32), ("w", C.c_int32), ("h", C.c_int32),
                ("npts", C.c_int32), ("first_px", C.c_int32), ("reserved", C.c_int32),
                ("sum_x", C.c_int64), ("sum_y", C.c_int64),
                ("area", C.c_double), ("m00", C.c_double), ("m10", C.c_double), ("m01", C.c_double)]


_lib = None


def load():
    global _lib
    if _lib is None:
        src = os.path.join(_HERE, "abub_ref.c")
        if not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
            subprocess.check_call(["make", "-s", "-C", _HERE])
        _lib = C.CDLL(_LIB)
        _lib.ref_otsu.restype = C.c_int
        _lib.ref_otsu.argtypes = [C.c_void_p, C.c_size_t]
        _lib.ref_tozero_otsu_binary.restype = C.c_int
        _lib.ref_tozero_otsu_binary.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
        _lib.ref_find_contours.restype = C.c_int
        _lib.ref_find_contours.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(RefBlob), C.c_int, C.c_void_p,
                                           C.c_int, C.POINTER(C.c_int)]
        _lib.ref_process_frame.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        _lib.ref_hist256.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
        _lib.ref_post_over.argtypes = [C.c_void_p] * 3 + [C.c_int, C.c_int, C.c_void_p]
        _lib.ref_welford.argtypes = [C.c_void_p, C.c_int, C.c_size_t, C.c_void_p, C.c_void_p]
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def process_frame(a, b, sigma):
    a, b, sigma = (np.ascontiguousarray(v, np.uint8) for v in (a, b, sigma))
    out = np.empty_like(a)
    load().ref_process_frame(_p(a), _p(b), _p(sigma), a.shape[0], a.shape[1], _p(out))
    return out


def hist256(img):
    img = np.ascontiguousarray(img, np.uint8)
    h = np.zeros(256, np.uint32)
    load().ref_hist256(_p(img), img.size, _p(h))
    return h


def post_over(frame, avg, sigma):
    frame, avg, sigma = (np.ascontiguousarray(v, np.uint8) for v in (frame, avg, sigma))
    out = np.empty_like(frame)
    load().ref_post_over(_p(frame), _p(avg), _p(sigma), frame.shape[0], frame.shape[1], _p(out))
    return out


def otsu(hist, total):
    h = np.ascontiguousarray(hist, np.uint32)
    return load().ref_otsu(_p(h), int(total))


def tozero_otsu_binary(src, tozero):
    src = np.ascontiguousarray(src, np.uint8)
    out = np.empty_like(src)
    t = load().ref_tozero_otsu_binary(_p(src), src.size, int(tozero), _p(out))
    return out, t


def find_contours(binary, cap=65536, pts_cap=1 << 20):
    """Returns (list of RefBlob, list of [n,2] int32 vertex arrays) in findContours order."""
    b = np.ascontiguousarray(binary, np.uint8)
    blobs = (RefBlob * cap)()
    pts = np.zeros((pts_cap, 2), np.int32)
    ntot = C.c_int()
    n = load().ref_find_contours(_p(b), b.shape[0], b.shape[1], blobs, cap, _p(pts), pts_cap, C.byref(ntot))
    assert 0 <= n <= cap and ntot.value <= pts_cap
    polys, off = [], 0
    for i in range(n):
        polys.append(pts[off:off + blobs[i].npts].copy())
        off += blobs[i].npts
    return [blobs[i] for i in range(n)], polys


def welford(frames):
    fr = [np.ascontiguousarray(f, np.uint8) for f in frames]
    n = len(fr)
    tab = (C.c_void_p * n)(*[f.ctypes.data for f in fr])
    mean = np.empty_like(fr[0])
    sig = np.empty_like(fr[0])
    load().ref_welford(tab, n, fr[0].size, _p(mean), _p(sig))
    return mean, sig
