"""ctypes front end of oracle/oracle_c.c  --  TEST INFRASTRUCTURE, NOT PRODUCT (see oracle/oracle.py)."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import build_oracle

_lib = None


def lib():
    global _lib
    if _lib is None:
        path = build_oracle.OUT
        if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(build_oracle.SRC):
            path = build_oracle.build()
        L = C.CDLL(path)
        vp, ci = C.c_void_p, C.c_int
        L.oc_init_maps.argtypes = [vp, vp, ci, vp, vp, ci, ci, ci, vp, vp]
        L.oc_remap.argtypes = [vp, ci, ci, vp, vp, ci, ci, vp]
        L.oc_preprocess.argtypes = [ci, vp, ci, ci, vp]
        L.oc_preprocess_f16.argtypes = [ci, vp, ci, ci, vp]
        L.oc_reproject.argtypes = [ci, vp, vp, vp, ci, ci, vp]
        L.oc_compact.argtypes = [vp, vp, C.c_long, vp]
        L.oc_compact.restype = C.c_long
        L.oc_heatmap.argtypes = [vp, ci, ci, vp, vp]
        L.oc_resize.argtypes = [vp, ci, ci, vp, ci, ci]
        L.oc_max_threads.restype = ci
        L.oc_set_threads.argtypes = [ci]
        _lib = L
    return _lib


def _p(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


def set_threads(n: int):
    lib().oc_set_threads(int(n))


def max_threads() -> int:
    return lib().oc_max_threads()


def init_maps(K, D, R, P, size):
    W, H = int(size[0]), int(size[1])
    K, D, R, P = _c(K, np.float64), _c(D, np.float64).ravel(), _c(R, np.float64), _c(P, np.float64)
    m1 = np.empty((H, W, 2), np.int16)
    m2 = np.empty((H, W), np.uint16)
    lib().oc_init_maps(_p(K), _p(D), D.size, _p(R), _p(P), P.shape[1], W, H, _p(m1), _p(m2))
    return m1, m2


def remap(src, map1, map2):
    src = _c(src, np.uint8)
    H, W = map2.shape
    dst = np.empty((H, W, 3), np.uint8)
    lib().oc_remap(_p(src), src.shape[0], src.shape[1], _p(_c(map1, np.int16)), _p(_c(map2, np.uint16)), H, W, _p(dst))
    return dst


def preprocess_eye(model, img, dtype=np.float32):
    img = _c(img, np.uint8)
    H, W = img.shape[:2]
    if np.dtype(dtype) == np.float16:                      # extension a10: rounded once to half
        out = np.empty((3, H, W), np.float16)
        lib().oc_preprocess_f16(int(model), _p(img), H, W, _p(out))
        return out
    out = np.empty((3, H, W), np.float32)
    lib().oc_preprocess(int(model), _p(img), H, W, _p(out))
    return out


def reproject(model, disparity, left, Q):
    d = np.array(disparity, np.float32, copy=True, order="C")
    left = _c(left, np.uint8)
    H, W = d.shape
    q = _c(np.asarray(Q, np.float64).reshape(16).astype(np.float32), np.float32)
    pts = np.empty((H * W, 6), np.float32)
    lib().oc_reproject(int(model), _p(d), _p(left), _p(q), H, W, _p(pts))
    return pts, d


def compact(disparity_used, dense):
    d = _c(disparity_used, np.float32).ravel()
    dense = _c(dense, np.float32)
    out = np.empty_like(dense)
    k = lib().oc_compact(_p(d), _p(dense), d.size, _p(out))
    return out[:k].copy(), int(k)


def heatmap(disp, jet):
    disp = _c(disp, np.float32)
    jet = _c(jet, np.uint8)
    H, W = disp.shape
    out = np.empty((H, W, 3), np.uint8)
    lib().oc_heatmap(_p(disp), H, W, _p(jet), _p(out))
    return out


def resize(src, size):
    src = _c(src, np.uint8)
    w, h = int(size[0]), int(size[1])
    dst = np.empty((h, w, 3), np.uint8)
    lib().oc_resize(_p(src), src.shape[0], src.shape[1], _p(dst), h, w)
    return dst
