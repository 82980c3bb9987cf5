"""Python face of the C ABI in include/stereo_geom.h: one thin method per sa_* entry point.

Host data are numpy arrays; device data are anything with a ``data_ptr()`` (torch CUDA tensors) or a raw integer
address.  Nothing here computes: every method forwards to libstereo_geom.so and raises ``StereoGeomError`` on a
negative status.  The reference-named entry points live in ``stereoalgorithms_b200.compat``.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib as L
from ._lib import (COMPACT, DENSE, DEVICE, F16, F32, HOST, LEFT, RIGHT, MODEL_CRESTEREO, MODEL_FASTACVNET_PLUS,  # noqa: F401
                   MODEL_HITNET, MODEL_RAFTSTEREO)

MODELS = {"RAFTStereo": MODEL_RAFTSTEREO, "HitNet": MODEL_HITNET, "CREStereo": MODEL_CRESTEREO,
          "FastACVNet_plus": MODEL_FASTACVNET_PLUS}


class StereoGeomError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"[{code}] {msg}")
        self.code = code


def _ptr(x):
    """(address, is_device) of a numpy array, a torch tensor, or an int."""
    if x is None:
        return None, False
    if isinstance(x, np.ndarray):
        if not x.flags["C_CONTIGUOUS"]:
            raise ValueError("array must be C-contiguous")
        return x.ctypes.data, False
    if hasattr(x, "data_ptr"):
        return int(x.data_ptr()), bool(getattr(x, "is_cuda", False))
    return int(x), True


class StereoGeometry:
    """sa_handle wrapper.  ``with StereoGeometry(...) as g:`` destroys the handle on exit."""

    def __init__(self, width: int, height: int, model="FastACVNet_plus", dtype=F32, max_batch: int = 1, gpu_id: int = 0,
                 net_size=None, flags: int = 0):
        self._lib = L.load()
        self.model = MODELS[model] if isinstance(model, str) else int(model)
        self.W, self.H = int(width), int(height)
        self.w, self.h = (int(net_size[0]), int(net_size[1])) if net_size else (self.W, self.H)
        self.dtype = dtype
        self.max_batch = int(max_batch)
        cfg = L.sa_config(C.sizeof(L.sa_config), gpu_id, self.W, self.H, self.w if net_size else 0, self.h if net_size else 0,
                          self.model, dtype, self.max_batch, flags)
        h = C.c_void_p()
        rc = self._lib.sa_create(C.byref(cfg), C.byref(h))
        if rc != L.SA_OK:
            raise StereoGeomError(rc, (self._lib.sa_last_error(None) or b"").decode())
        self._h = h
        self.planes = self._lib.sa_disparity_planes(self._h)
        self._engine_ref = None

    # -- plumbing ---------------------------------------------------------------------------------------------
    def _check(self, rc: int):
        if rc != L.SA_OK:
            raise StereoGeomError(rc, (self._lib.sa_last_error(self._h) or b"").decode())

    def close(self):
        if getattr(self, "_h", None):
            self._lib.sa_destroy(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def handle(self):
        return self._h

    @property
    def np_dtype(self):
        return np.float16 if self.dtype == F16 else np.float32

    # -- calibration ------------------------------------------------------------------------------------------
    def load_calibration_yaml(self, path: str):
        self._check(self._lib.sa_load_calibration_yaml(self._h, path.encode()))

    def set_calibration(self, calib: dict):
        """calib: the 11 matrices of a StereoCalibration.yml as float64 arrays (keys as in the file)."""
        a = {k: np.ascontiguousarray(np.asarray(v, np.float64)) for k, v in calib.items()}
        dl, dr = a["distCoeffs_left"].ravel(), a["distCoeffs_right"].ravel()
        P1 = a["P1"] if a["P1"].shape == (3, 4) else np.ascontiguousarray(np.hstack([a["P1"], np.zeros((3, 1))]))
        P2 = a["P2"] if a["P2"].shape == (3, 4) else np.ascontiguousarray(np.hstack([a["P2"], np.zeros((3, 1))]))
        q = a.get("Q")
        self._check(self._lib.sa_set_calibration(self._h, a["intrinsic_left"].ctypes.data, dl.ctypes.data, dl.size, a["R_L"].ctypes.data,
                                                 P1.ctypes.data, a["intrinsic_right"].ctypes.data, dr.ctypes.data, dr.size,
                                                 a["R_R"].ctypes.data, P2.ctypes.data, q.ctypes.data if q is not None else None))

    def set_Q(self, Q):
        q = np.ascontiguousarray(np.asarray(Q, np.float64).reshape(16))
        self._check(self._lib.sa_set_Q(self._h, q.ctypes.data))

    def get_Q(self):
        q = np.empty(16, np.float64)
        self._check(self._lib.sa_get_Q(self._h, q.ctypes.data))
        return q.reshape(4, 4)

    def set_maps(self, eye: int, map1: np.ndarray, map2: np.ndarray):
        m1 = np.ascontiguousarray(map1, np.int16)
        m2 = np.ascontiguousarray(map2, np.uint16)
        assert m1.shape == (self.H, self.W, 2) and m2.shape == (self.H, self.W)
        self._check(self._lib.sa_set_maps(self._h, eye, m1.ctypes.data, m2.ctypes.data))

    def get_maps(self, eye: int):
        m1 = np.empty((self.H, self.W, 2), np.int16)
        m2 = np.empty((self.H, self.W), np.uint16)
        self._check(self._lib.sa_get_maps(self._h, eye, m1.ctypes.data, m2.ctypes.data))
        return m1, m2

    # -- stages -------------------------------------------------------------------------------------------------
    def rectify_preprocess(self, left, right, n: int | None = None, rectify: bool = True, pitch: int = 0, stream=None):
        """left/right: uint8 [n,H,W,3] (numpy = host, torch CUDA tensor = device)."""
        pl, dev_l = _ptr(left)
        pr, dev_r = _ptr(right)
        if dev_l != dev_r:
            raise ValueError("left and right must both be host or both be device")
        if n is None:
            n = int(left.shape[0]) if getattr(left, "ndim", 3) == 4 else 1
        self._check(self._lib.sa_rectify_preprocess(self._h, n, pl, pr, pitch, DEVICE if dev_l else HOST, 1 if rectify else 0, stream))
        return n

    def rectify_preprocess_side_by_side(self, frames, rectify: bool = True, stream=None):
        """frames: uint8 [n, H, 2W, 3], left half | right half - the 1280x480 capture the reference's 3-argument
        RectifyImage overload splits with two clone()s (FastACVNet_plus/src/FastACVNet_plus_Algorithm.cpp:116-131).
        Zero copy here: the two eyes are the same buffer at a byte offset of 3*W with a row pitch of 6*W."""
        p, dev = _ptr(frames)
        n = int(frames.shape[0])
        if tuple(frames.shape[1:]) != (self.H, 2 * self.W, 3):
            raise ValueError(f"expected [n, {self.H}, {2 * self.W}, 3], got {tuple(frames.shape)}")
        # consecutive frames are 6*W*H bytes apart = H rows of the doubled pitch, which is what the C ABI assumes
        self._check(self._lib.sa_rectify_preprocess(self._h, n, p, p + 3 * self.W, 6 * self.W, DEVICE if dev else HOST, 1 if rectify else 0, stream))
        return n

    @staticmethod
    def _jpeg_args(jpegs):
        bufs = [bytes(j) for j in jpegs]                       # (kept alive by the caller's frame: ctypes only borrows them)
        ptrs = (C.c_void_p * len(bufs))(*[C.cast(C.c_char_p(b), C.c_void_p) for b in bufs])
        sizes = (C.c_size_t * len(bufs))(*[len(b) for b in bufs])
        return bufs, ptrs, sizes

    def decode_capture_jpeg(self, jpegs, stream=None) -> int:
        """jpegs: n compressed side-by-side capture frames (bytes; 2W x H each) - what the reference's MJPG camera delivers
        (usb_test.py:7-11) - decoded on the GPU into the handle's capture buffer.  -> n"""
        bufs, ptrs, sizes = self._jpeg_args(jpegs)
        self._check(self._lib.sa_decode_capture_jpeg(self._h, len(bufs), ptrs, sizes, stream))
        return len(bufs)

    def rectify_preprocess_jpeg(self, jpegs, rectify: bool = True, stream=None) -> int:
        """decode_capture_jpeg + stage A on the two halves of the decoded frames, in place (no split, no copy)."""
        bufs, ptrs, sizes = self._jpeg_args(jpegs)
        self._check(self._lib.sa_rectify_preprocess_jpeg(self._h, len(bufs), ptrs, sizes, 1 if rectify else 0, stream))
        return len(bufs)

    def encode_jpeg(self, image, quality: int = 95, stream=None) -> bytes:
        """image: uint8 [H, W] (grey) or [H, W, 3] (BGR), numpy or a CUDA tensor -> the bytes of a baseline JPEG (what the demos'
        cv::imwrite("disparity.jpg" / "heatmap.jpg") store, FastACVNet_plus/test/main.cpp:55,59; encoded on the GPU)."""
        p, dev = _ptr(image)
        hh, ww = int(image.shape[0]), int(image.shape[1])
        ch = 1 if len(image.shape) == 2 else int(image.shape[2])
        need = C.c_size_t(0)
        where = DEVICE if dev else HOST
        self._check(self._lib.sa_encode_jpeg(self._h, p, where, ww, hh, ch, 0, quality, None, 0, C.byref(need), stream))
        buf = (C.c_ubyte * need.value)()
        self._check(self._lib.sa_encode_jpeg(self._h, p, where, ww, hh, ch, 0, quality, buf, need.value, C.byref(need), stream))
        return bytes(buf[: need.value])

    def write_jpeg(self, path: str, image, quality: int = 95):
        p, dev = _ptr(image)
        ch = 1 if len(image.shape) == 2 else int(image.shape[2])
        self._check(self._lib.sa_write_jpeg(self._h, path.encode(), p, DEVICE if dev else HOST, int(image.shape[1]), int(image.shape[0]), ch, 0, quality))

    def decode_jpeg(self, data: bytes) -> np.ndarray:
        """One JPEG stream of any size -> uint8 [H, W, 3] BGR (decoded on the GPU)."""
        b = bytes(data)
        w, hh = C.c_int(0), C.c_int(0)
        self._check(self._lib.sa_decode_jpeg(self._h, b, len(b), None, HOST, 0, C.byref(w), C.byref(hh), None))
        out = np.empty((hh.value, w.value, 3), np.uint8)
        self._check(self._lib.sa_decode_jpeg(self._h, b, len(b), out.ctypes.data, HOST, 0, C.byref(w), C.byref(hh), None))
        return out

    def capture_ptr(self) -> int:
        return int(self._lib.sa_capture_buffer(self._h) or 0)

    def capture(self, n: int, stream=None) -> np.ndarray:
        """The first n decoded capture frames, uint8 [n, H, 2W, 3] (host copy; synchronises)."""
        out = np.empty((n, self.H, 2 * self.W, 3), np.uint8)
        self._check(self._lib.sa_copy_capture(self._h, n, out.ctypes.data, HOST, stream))
        return out

    def input_tensor_ptr(self, eye: int = LEFT) -> int:
        return int(self._lib.sa_input_tensor(self._h, eye))

    def disparity_ptr(self) -> int:
        return int(self._lib.sa_disparity_buffer(self._h))

    def pointcloud_ptr(self) -> int:
        return int(self._lib.sa_pointcloud_buffer(self._h) or 0)

    def stub_disparity(self, n: int, seed: int = 1, stream=None):
        self._check(self._lib.sa_stub_disparity(self._h, n, seed, stream))

    def set_disparity(self, disparity, n: int | None = None, stream=None):
        """disparity: float32 [n,planes,h,w] network output (host numpy or device tensor)."""
        if isinstance(disparity, np.ndarray):
            disparity = np.ascontiguousarray(disparity, np.float32)
        p, dev = _ptr(disparity)
        if n is None:
            n = int(np.prod(disparity.shape)) // (self.planes * self.w * self.h)
        self._check(self._lib.sa_set_disparity(self._h, n, p, DEVICE if dev else HOST, stream))
        self._keep = disparity

    def reproject(self, n: int, mode: int = DENSE, out=None, counts=None, to_host: bool = True, stream=None):
        """Returns (points, counts).  to_host: numpy [n,h*w,6] float32 (+ int32 [n]); compact frames keep their
        fixed h*w-row slot, rows beyond counts[i] are left untouched (zeros here).  With to_host=False results stay in
        `out`/`counts` if given (device) or in the handle's internal buffer."""
        npix = self.w * self.h
        if to_host:
            pts = np.zeros((n, npix, 6), np.float32) if out is None else out
            cnt = np.zeros(n, np.int32) if counts is None else counts
            self._check(self._lib.sa_reproject(self._h, n, mode, pts.ctypes.data, cnt.ctypes.data, HOST, stream))
            return pts, cnt
        po, _ = _ptr(out)
        pc, _ = _ptr(counts)
        self._check(self._lib.sa_reproject(self._h, n, mode, po, pc, DEVICE, stream))
        return out, counts

    def heatmap(self, n: int, out=None, to_host: bool = True, stream=None):
        if to_host:
            hm = np.empty((n, self.h, self.w, 3), np.uint8) if out is None else out
            self._check(self._lib.sa_heatmap(self._h, n, hm.ctypes.data, HOST, stream))
            return hm
        po, _ = _ptr(out)
        self._check(self._lib.sa_heatmap(self._h, n, po, DEVICE, stream))
        return out

    def disparity_u8(self, n: int, out=None, to_host: bool = True, stream=None):
        """The disparity as cv::imwrite("disparity.jpg", disparity) sees it (FastACVNet_plus/test/main.cpp:55):
        convertTo(CV_8U) of the float image -> uint8 [n, h, w]."""
        if to_host:
            img = np.empty((n, self.h, self.w), np.uint8) if out is None else out
            self._check(self._lib.sa_disparity_u8(self._h, n, img.ctypes.data, HOST, stream))
            return img
        po, _ = _ptr(out)
        self._check(self._lib.sa_disparity_u8(self._h, n, po, DEVICE, stream))
        return out

    # -- read-backs ---------------------------------------------------------------------------------------------
    def rectified(self, n: int, eye: int) -> np.ndarray:
        out = np.empty((n, self.H, self.W, 3), np.uint8)
        self._check(self._lib.sa_copy_rectified(self._h, n, eye, out.ctypes.data, HOST, None))
        return out

    def net_image(self, n: int, eye: int) -> np.ndarray:
        out = np.empty((n, self.h, self.w, 3), np.uint8)
        self._check(self._lib.sa_copy_net_image(self._h, n, eye, out.ctypes.data, HOST, None))
        return out

    def tensor(self, n: int, eye: int = LEFT) -> np.ndarray:
        ch = 6 if self.model == MODEL_HITNET else 3
        out = np.empty((n, ch, self.h, self.w), self.np_dtype)
        self._check(self._lib.sa_copy_tensor(self._h, n, eye, out.ctypes.data, HOST, None))
        return out

    def disparity(self, n: int) -> np.ndarray:
        out = np.empty((n, self.h, self.w), np.float32)
        self._check(self._lib.sa_copy_disparity(self._h, n, out.ctypes.data, HOST, None))
        return out

    def synchronize(self):
        self._check(self._lib.sa_synchronize(self._h))

    @property
    def stream(self) -> int:
        """The handle's own compute stream (sa_get_stream) as an integer cudaStream_t."""
        return int(self._lib.sa_get_stream(self._h) or 0)

    def download(self, n: int, pointcloud=None, disparity=None, rect_left=None, rect_right=None, heatmap=None, stream=None):
        """sa_download: every requested read-back (host numpy arrays) queued on one stream, one synchronisation."""
        def img(a):          # [n, H, W, 3] uint8 whose rows may be padded (a view of a wider array), frames H rows apart
            if a is None:
                return None, 0
            if a.strides[-1] != 1 or a.strides[-2] != 3 or (a.ndim == 4 and a.strides[0] != a.strides[1] * self.H):
                raise ValueError("rectified image destination: packed pixels, frames H rows apart")
            return a.ctypes.data, int(a.strides[-3])
        (pl, sl), (pr, sr) = img(rect_left), img(rect_right)
        ptrs = [_ptr(x)[0] for x in (pointcloud, disparity, heatmap)]
        self._check(self._lib.sa_download(self._h, n, ptrs[0], ptrs[1], pl, sl, pr, sr, ptrs[2], stream))

    # -- whole path with host buffers ----------------------------------------------------------------------------
    def set_engine(self, fn):
        """fn(stream, first_frame, n, left_tensor_ptr, right_tensor_ptr, disparity_ptr) -> int, or None."""
        if fn is None:
            self._engine_ref = None
            self._check(self._lib.sa_set_engine(self._h, L.ENGINE_FN(), None))
            return
        cb = L.ENGINE_FN(lambda user, stream, f0, n, lt, rt, d: int(fn(stream, f0, n, lt, rt, d) or 0))
        self._engine_ref = cb
        self._check(self._lib.sa_set_engine(self._h, cb, None))

    def run_batch(self, left, right, n: int, rectify: bool = True, stub_seed: int = 0, pointcloud=None, disparity=None,
                  rect_left=None, rect_right=None, heatmap=None):
        """All buffers are HOST memory (numpy arrays or pinned addresses)."""
        ptrs = [_ptr(x)[0] for x in (left, right, pointcloud, disparity, rect_left, rect_right, heatmap)]
        self._check(self._lib.sa_run_batch(self._h, n, ptrs[0], ptrs[1], 1 if rectify else 0, stub_seed, *ptrs[2:]))

    # -- observability -------------------------------------------------------------------------------------------
    def stage_time(self, stage: int):
        ms, cnt = C.c_double(), C.c_long()
        self._check(self._lib.sa_stage_time(self._h, stage, C.byref(ms), C.byref(cnt)))
        return ms.value, cnt.value

    def reset_timing(self):
        self._check(self._lib.sa_reset_timing(self._h))

    @property
    def launch_count(self) -> int:
        return int(self._lib.sa_launch_count(self._h))

    def set_option(self, key: str, value: int):
        self._check(self._lib.sa_set_option(self._h, key.encode(), int(value)))


def run_batch_multi(geoms, left, right, n: int, rectify: bool = True, stub_seed: int = 0, pointcloud=None, disparity=None,
                    rect_left=None, rect_right=None, heatmap=None):
    """sa_run_batch_multi: one host batch of n pairs cut frame-wise over the handles in ``geoms`` (one per GPU), each shard on
    its own host thread.  All buffers are HOST memory with the layouts of ``StereoGeometry.run_batch``."""
    lib = L.load()
    arr = (C.c_void_p * len(geoms))(*[g.handle for g in geoms])
    ptrs = [_ptr(x)[0] for x in (left, right, pointcloud, disparity, rect_left, rect_right, heatmap)]
    failed = C.c_int(-1)
    rc = lib.sa_run_batch_multi(arr, len(geoms), n, ptrs[0], ptrs[1], 1 if rectify else 0, stub_seed, *ptrs[2:], C.byref(failed))
    if rc != L.SA_OK:
        msg = (lib.sa_last_error(geoms[failed.value].handle) or b"").decode() if failed.value >= 0 else "invalid argument"
        raise StereoGeomError(rc, f"shard {failed.value}: {msg}")


class PinnedArray:
    """numpy view over cudaMallocHost memory (sa_host_alloc).  Keep the object alive while ``.array`` is in use."""

    def __init__(self, shape, dtype):
        self._lib = L.load()
        nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = C.c_void_p()
        rc = self._lib.sa_host_alloc(C.byref(p), nbytes)
        if rc != L.SA_OK:
            raise StereoGeomError(rc, (self._lib.sa_last_error(None) or b"").decode())
        self._addr = p.value
        self._buf = (C.c_uint8 * max(nbytes, 1)).from_address(p.value)
        self.array = np.frombuffer(self._buf, dtype=np.uint8, count=nbytes).view(dtype).reshape(shape)

    def free(self):
        if getattr(self, "_addr", None):
            self.array = None
            self._buf = None
            self._lib.sa_host_free(self._addr)
            self._addr = None

    __del__ = free


def stereo_rectify(K_l, D_l, K_r, D_r, size, R, T, flags: int = 1024, alpha: float = 0.0, new_size=(0, 0)):
    """cv::stereoRectify as the reference's calibration tool calls it (Stereo_Calibration/Stereo_Calibration.cpp:162:
    CALIB_ZERO_DISPARITY = 1024, alpha 0, new size = image size) -> R_L, R_R (3x3), P1, P2 (3x4), Q (4x4), roi_l, roi_r
    (x, y, w, h).  Host code (sa_stereo_rectify); needs no GPU."""
    lib = L.load()
    a = lambda x: np.ascontiguousarray(np.asarray(x, np.float64).ravel())
    K_l, D_l, K_r, D_r, R, T = (a(v) for v in (K_l, D_l, K_r, D_r, R, T))
    if K_l.size != 9 or K_r.size != 9 or R.size != 9 or T.size != 3:
        raise ValueError("K_l, K_r, R must have 9 elements and T 3")
    R1, R2, P1, P2, Q = np.zeros((3, 3)), np.zeros((3, 3)), np.zeros((3, 4)), np.zeros((3, 4)), np.zeros((4, 4))
    r1, r2 = np.zeros(4, np.int32), np.zeros(4, np.int32)
    rc = lib.sa_stereo_rectify(K_l.ctypes.data, D_l.ctypes.data, D_l.size, K_r.ctypes.data, D_r.ctypes.data, D_r.size, int(size[0]), int(size[1]),
                               R.ctypes.data, T.ctypes.data, int(flags), float(alpha), int(new_size[0]), int(new_size[1]), R1.ctypes.data, R2.ctypes.data,
                               P1.ctypes.data, P2.ctypes.data, Q.ctypes.data, r1.ctypes.data, r2.ctypes.data)
    if rc != 0:
        raise StereoGeomError(rc, "sa_stereo_rectify: invalid argument")
    return R1, R2, P1, P2, Q, tuple(int(v) for v in r1), tuple(int(v) for v in r2)
