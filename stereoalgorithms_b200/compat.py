"""The reference's per-model entry points, by their own names, on top of the C ABI (sa_ref_* in
include/stereo_algorithms_compat.h).

The reference ships one shared library per model; each exports ``Initialize``, ``Run<Model>[_RectifyImage]``,
``Version`` and ``Release`` (e.g. FastACVNet_plus/include/FastACVNet_plus_Algorithm.h:23-43).  Here every model is a
namespace object with those same functions::

    from stereoalgorithms_b200.compat import FastACVNet_plus as F
    p = F.Initialize("model.onnx", 0, "StereoCalibration.yml")
    pointcloud = np.empty(640 * 480 * 6, np.float32)           # float* pointcloud = new float[W*H*6]
    disparity = Mat()                                          # cv::Mat disparity;
    rc = F.RunFastACVNet_plus_RectifyImage(p, left, right, pointcloud, disparity)
    F.Release(p)

Argument meaning and error behaviour follow the reference: ``left``/``right`` are uint8 HxWx3 BGR arrays and are
overwritten with the rectified images by the rectifying calls; ``pointcloud`` is caller-allocated and receives H*W
rows of x y z r g b; ``disparity.array`` is replaced by a fresh HxW float32 array; the return value is 0, or -1 after
a failed ``Initialize`` or for empty images; ``Initialize`` always returns a handle.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib as L


class Mat:
    """Stand-in for the ``cv::Mat& disparity`` out-parameter: ``array`` is re-assigned by every Run call."""

    def __init__(self, array=None):
        self.array = array

    def empty(self) -> bool:
        return self.array is None or self.array.size == 0


def _image(a):
    if a is None or not isinstance(a, np.ndarray) or a.size == 0:
        return L.sa_image(None, 0, 0, 0)
    if a.dtype != np.uint8 or a.ndim != 3 or a.shape[2] != 3 or a.strides[2] != 1 or a.strides[1] != 3:
        raise TypeError("images must be uint8 HxWx3 arrays with packed pixels (CV_8UC3)")
    return L.sa_image(a.ctypes.data, a.shape[0], a.shape[1], a.strides[0])


class _Model:
    def __init__(self, model_id: int, run_name: str, has_plain_run: bool):
        self._id = model_id
        self._lib = None
        # RAFTStereo / HitNet export one Run that always rectifies (RAFTStereo/src/RAFTStereoAlgorithm.cpp:57-72,
        # HitNet/src/HitNetAlgorithm.cpp:40-50); CREStereo / FastACVNet_plus export Run<Model> (no rectification) and
        # Run<Model>_RectifyImage (CREStereo/include/CREStereoAlgorithm.h:29,34).
        if has_plain_run:
            setattr(self, run_name, lambda p, l, r, pc, d: self._run(p, 0, l, r, pc, d))
            setattr(self, run_name + "_RectifyImage", lambda p, l, r, pc, d: self._run(p, 1, l, r, pc, d))
        else:
            setattr(self, run_name, lambda p, l, r, pc, d: self._run(p, 1, l, r, pc, d))

    def _l(self):
        if self._lib is None:
            self._lib = L.load()
        return self._lib

    def Initialize(self, model_path, gpu_id, calibration_path):
        enc = lambda s: s.encode() if isinstance(s, str) else s
        return self._l().sa_ref_Initialize(self._id, enc(model_path), int(gpu_id), enc(calibration_path))

    def _run(self, p, rectify, left, right, pointcloud, disparity):
        li, ri = _image(left), _image(right)
        rows, cols = li.rows, li.cols
        if not isinstance(pointcloud, np.ndarray) or pointcloud.dtype != np.float32 or not pointcloud.flags["C_CONTIGUOUS"]:
            raise TypeError("pointcloud must be a contiguous float32 array of H*W*6 elements")
        if rows > 0 and cols > 0 and pointcloud.size < rows * cols * 6:
            raise ValueError("pointcloud is too small for H*W*6 floats")
        d = np.empty((max(rows, 0), max(cols, 0)), np.float32)
        rc = self._l().sa_ref_Run(p, rectify, C.byref(li), C.byref(ri), pointcloud.ctypes.data, d.ctypes.data if d.size else None)
        if rc == 0 and disparity is not None:
            disparity.array = d          # DisparityMap = DisparityImage.clone() (TRTFastACVNet_plus.cpp:167-168)
        return rc

    def Version(self, p) -> str:
        return self._l().sa_ref_Version(p).decode()

    def Release(self, p) -> int:
        return self._l().sa_ref_Release(p)

    # not in the reference: diagnostics and the network slot
    def LastError(self, p) -> str:
        return self._l().sa_ref_LastError(p).decode()

    def SetStubSeed(self, p, seed: int) -> int:
        return self._l().sa_ref_SetStubSeed(p, int(seed))

    def SetEngine(self, p, fn) -> int:
        """The network slot (sa_ref_SetEngine): fn(stream, first_frame, n, left_tensor_ptr, right_tensor_ptr, disparity_ptr)
        -> int is called between the two halves of every Run with the handle's stream (an integer cudaStream_t, never 0):
        enqueue the network on that stream, do not synchronise.  None removes the callback."""
        if not hasattr(self, "_engines"):
            self._engines = {}
        cb = L.ENGINE_FN(lambda user, stream, f0, n, lt, rt, d: int(fn(stream, f0, n, lt, rt, d) or 0)) if fn else L.ENGINE_FN()
        self._engines[p] = cb          # keep the trampoline alive as long as the handle may call it
        return self._l().sa_ref_SetEngine(p, cb, None)

    def SetHostRegister(self, p, on: bool) -> int:
        """Page-locking of a reused pointcloud buffer (sa_ref_SetHostRegister); off by default."""
        return self._l().sa_ref_SetHostRegister(p, 1 if on else 0)

    def Geometry(self, p):
        return self._l().sa_ref_Geometry(p)


RAFTStereo = _Model(L.MODEL_RAFTSTEREO, "RunRAFTStereo", False)
HitNet = _Model(L.MODEL_HITNET, "RunHitNet", False)
CREStereo = _Model(L.MODEL_CRESTEREO, "RunCREStereo", True)
FastACVNet_plus = _Model(L.MODEL_FASTACVNET_PLUS, "RunFastACVNet_plus", True)


def heatmap(p, disparity: Mat) -> np.ndarray:
    """heatmap(cv::Mat& disparity) of the demos (FastACVNet_plus/test/main.cpp:7-24) for the disparity of the last
    Run on handle ``p``: BGR uint8 HxWx3."""
    lib = L.load()
    h, w = disparity.array.shape
    out = np.empty((h, w, 3), np.uint8)
    if lib.sa_ref_Heatmap(p, out.ctypes.data) != 0:
        raise RuntimeError(lib.sa_ref_LastError(p).decode())
    return out


def write_pointcloud_txt(path: str, pointcloud: np.ndarray) -> None:
    """The demo's pointcloud.txt (FastACVNet_plus/test/main.cpp:62-69)."""
    pts = np.ascontiguousarray(pointcloud, np.float32).reshape(-1, 6)
    rc = L.load().sa_write_pointcloud_txt(path.encode(), pts.ctypes.data, pts.shape[0])
    if rc != 0:
        raise OSError(f"sa_write_pointcloud_txt({path}) failed with {rc}")


def write_pointcloud_ply(path: str, pointcloud: np.ndarray, binary: bool = True, skip_non_finite: bool = True) -> None:
    """x y z r g b points as a PLY file (binary little-endian by default)."""
    pts = np.ascontiguousarray(pointcloud, np.float32).reshape(-1, 6)
    rc = L.load().sa_write_pointcloud_ply(path.encode(), pts.ctypes.data, pts.shape[0], 1 if binary else 0, 1 if skip_non_finite else 0)
    if rc != 0:
        raise OSError(f"sa_write_pointcloud_ply({path}) failed with {rc}")


def write_pointcloud_pcd(path: str, pointcloud: np.ndarray, skip_non_finite: bool = True) -> None:
    """x y z r g b points as a binary PCD v0.7 file (FIELDS x y z rgb)."""
    pts = np.ascontiguousarray(pointcloud, np.float32).reshape(-1, 6)
    rc = L.load().sa_write_pointcloud_pcd(path.encode(), pts.ctypes.data, pts.shape[0], 1 if skip_non_finite else 0)
    if rc != 0:
        raise OSError(f"sa_write_pointcloud_pcd({path}) failed with {rc}")


def write_pgm(path: str, grey: np.ndarray) -> None:
    """8-bit grey image as binary PGM - lossless stand-in for cv::imwrite("disparity.jpg", ...) of the demos."""
    g = np.ascontiguousarray(grey, np.uint8)
    rc = L.load().sa_write_pgm(path.encode(), g.ctypes.data, g.shape[1], g.shape[0], 0)
    if rc != 0:
        raise OSError(f"sa_write_pgm({path}) failed with {rc}")


def write_ppm(path: str, bgr: np.ndarray) -> None:
    """BGR uint8 HxWx3 image (rectified frame, heat map) as binary PPM (stored RGB)."""
    b = np.ascontiguousarray(bgr, np.uint8)
    rc = L.load().sa_write_ppm_bgr(path.encode(), b.ctypes.data, b.shape[1], b.shape[0], 0)
    if rc != 0:
        raise OSError(f"sa_write_ppm_bgr({path}) failed with {rc}")
