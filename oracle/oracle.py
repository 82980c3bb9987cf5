"""CPU oracle for the stereo geometry hot path  --  TEST INFRASTRUCTURE, NOT PRODUCT.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may import
this module.  The product (``stereoalgorithms_b200`` / ``libstereo_geom.so``) never does.

It restates, in numpy, the arithmetic of the reference's per-frame geometry path:

  stage                     reference call site                                   executes in
  ------------------------  ----------------------------------------------------  -----------------
  calibration YAML          FastACVNet_plus/src/FastACVNet_plus_Algorithm.cpp:98   OpenCV FileStorage
  rectification maps        FastACVNet_plus/src/FastACVNet_plus_Algorithm.cpp:139  OpenCV calib3d
  remap INTER_LINEAR        FastACVNet_plus/src/FastACVNet_plus_Algorithm.cpp:142  OpenCV imgproc
  preprocess (4 variants)   */src/*_preprocess.cu (kernel bodies)                  reference CUDA
  reproject (4 variants)    */src/*_preprocess.cu (reprojectImageTo3D kernels)     reference CUDA
  heatmap                   FastACVNet_plus/test/main.cpp:7-24                      OpenCV core/imgproc

OpenCV is a third-party dependency of the reference: it is NOT vendored under the reference
tree and NOT version-pinned (``find_package(OpenCV REQUIRED)``,
FastACVNet_plus/src/CMakeLists.txt:9).  The OpenCV stages are therefore restated from
OpenCV's published algorithm (calib3d ``initUndistortRectifyMap`` with ``CV_16SC2`` output,
imgproc ``remap`` fixed-point bilinear with 5 fractional bits and 15-bit weights, imgproc
8U ``resize`` INTER_LINEAR, ``applyColorMap(COLORMAP_JET)``) and PINNED against Python
``cv2`` 4.13.0 (opencv-python-headless 4.13.0.92) on the reference's own fixtures: see
``oracle/gen_golden.py`` (run in the build container, where /root/reference exists) and
``tests/test_oracle_golden.py``.  The reference itself ships no tests, golden vectors or
known-answer values for this path, so the pin is "cv2 executing the same
functions with the same arguments" plus the committed golden vectors it produced.

Everything here is deliberately simple, loop-free numpy; exactness over speed.
"""
from __future__ import annotations

import math
import os
import re

import numpy as np

# ----------------------------------------------------------------------------------------------
# model ids (shared vocabulary with include/stereo_geom.h; duplicated on purpose so the oracle
# does not import the product)
# ----------------------------------------------------------------------------------------------
RAFTSTEREO, HITNET, CRESTEREO, FASTACVNET_PLUS = 0, 1, 2, 3
MODEL_NAMES = {RAFTSTEREO: "RAFTStereo", HITNET: "HitNet", CRESTEREO: "CREStereo",
               FASTACVNET_PLUS: "FastACVNet_plus"}

INTER_BITS = 5
INTER_TAB_SIZE = 1 << INTER_BITS            # 32 sub-pixel positions per axis
INTER_REMAP_COEF_BITS = 15                  # weights sum to 1 << 15

CALIB_KEYS = ("intrinsic_left", "distCoeffs_left", "intrinsic_right", "distCoeffs_right",
              "R", "T", "R_L", "R_R", "P1", "P2", "Q")


# ----------------------------------------------------------------------------------------------
# a1. calibration file  (ReadObjectYml, FastACVNet_plus_Algorithm.cpp:98-114)
# ----------------------------------------------------------------------------------------------
def load_calibration_yaml(path: str) -> dict:
    """Parse the ``%YAML:1.0`` / ``!!opencv-matrix`` file cv::FileStorage writes
    (Stereo_Calibration/Stereo_Calibration.cpp:165-179) and ReadObjectYml reads.

    Returns {name: float64 ndarray (rows, cols)} for every ``!!opencv-matrix`` node found.
    Plain flow lists (validROIL/validROIR) and unknown keys are ignored, like the reference.
    """
    with open(path, "r") as f:
        text = f.read()
    out = {}
    # a node is `name: !!opencv-matrix` followed by indented rows/cols/dt/data lines
    pat = re.compile(r"^([A-Za-z_][A-Za-z0-9_]*)\s*:\s*!!opencv-matrix\s*$", re.M)
    heads = list(pat.finditer(text))
    for n, m in enumerate(heads):
        body = text[m.end(): heads[n + 1].start() if n + 1 < len(heads) else len(text)]
        rows = int(re.search(r"rows\s*:\s*(\d+)", body).group(1))
        cols = int(re.search(r"cols\s*:\s*(\d+)", body).group(1))
        dt = re.search(r"dt\s*:\s*\"?([0-9]*[a-z])\"?", body).group(1)
        data = re.search(r"data\s*:\s*\[(.*?)\]", body, re.S).group(1)
        vals = [float(tok) for tok in data.replace("\n", " ").split(",") if tok.strip()]
        if dt not in ("d", "f"):
            raise ValueError(f"{path}: node {m.group(1)}: unsupported dt '{dt}'")
        if len(vals) != rows * cols:
            raise ValueError(f"{path}: node {m.group(1)}: {len(vals)} values for {rows}x{cols}")
        out[m.group(1)] = np.asarray(vals, np.float64).reshape(rows, cols)
    return out


def write_calibration_yaml(path: str, calib: dict) -> None:
    """Write matrices in the layout cv::FileStorage produces (used to make fixtures)."""
    lines = ["%YAML:1.0", "---"]
    for k, m in calib.items():
        m = np.asarray(m, np.float64)
        m2 = m.reshape(m.shape[0], -1) if m.ndim == 2 else m.reshape(1, -1)
        lines += [f"{k}: !!opencv-matrix", f"   rows: {m2.shape[0]}", f"   cols: {m2.shape[1]}",
                  "   dt: d"]
        toks = [_fmt_yaml_double(v) for v in m2.ravel()]
        # wrap like OpenCV does: continuation lines are indented by 7 spaces
        cur, rows_txt = "   data: [ ", []
        for i, t in enumerate(toks):
            piece = t + (", " if i + 1 < len(toks) else " ]")
            if len(cur) + len(piece) > 76 and cur.strip():
                rows_txt.append(cur.rstrip())
                cur = "       "
            cur += piece
        rows_txt.append(cur)
        lines += rows_txt
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")


def _fmt_yaml_double(v: float) -> str:
    if v == int(v) and abs(v) < 1e15:
        return f"{int(v)}."
    return np.format_float_scientific(v, precision=16, unique=True, exp_digits=2)


def scale_calibration(calib: dict, s: float, crop_y: float = 0.0, crop_x: float = 0.0) -> dict:
    """Isotropic scale of K_l, K_r, P1, P2 rows 0-1 by ``s`` followed by a crop of
    (crop_x, crop_y) pixels off the principal points; Q rebuilt.
    """
    c = {k: np.array(v, np.float64, copy=True) for k, v in calib.items()}
    for k in ("intrinsic_left", "intrinsic_right", "P1", "P2"):
        c[k][0:2, :] *= s
        c[k][0, 2] -= crop_x
        c[k][1, 2] -= crop_y
    q = c["Q"]
    q[0, 3] = -c["P1"][0, 2]
    q[1, 3] = -c["P1"][1, 2]
    q[2, 3] = c["P1"][0, 0]
    # Q[3,2] (= -1/Tx) is unchanged by an isotropic scale; Q[3,3] = (cx1-cx2)/Tx scales with s
    q[3, 3] *= s
    return c


# ----------------------------------------------------------------------------------------------
# a2. cv::initUndistortRectifyMap(K, D, R, P, size, CV_16SC2)
#     call sites FastACVNet_plus_Algorithm.cpp:139-140 (and :126-127, Stereo_Calibration.cpp:244)
# ----------------------------------------------------------------------------------------------
def invert3x3(a: np.ndarray) -> np.ndarray:
    """Closed-form adjugate inverse in float64 (what cv::invert does for 3x3 DECOMP_LU)."""
    a = np.asarray(a, np.float64)
    d = (a[0, 0] * (a[1, 1] * a[2, 2] - a[1, 2] * a[2, 1])
         - a[0, 1] * (a[1, 0] * a[2, 2] - a[1, 2] * a[2, 0])
         + a[0, 2] * (a[1, 0] * a[2, 1] - a[1, 1] * a[2, 0]))
    if d == 0.0:
        return np.zeros((3, 3))
    d = 1.0 / d
    t = np.empty((3, 3))
    t[0, 0] = (a[1, 1] * a[2, 2] - a[1, 2] * a[2, 1]) * d
    t[0, 1] = (a[0, 2] * a[2, 1] - a[0, 1] * a[2, 2]) * d
    t[0, 2] = (a[0, 1] * a[1, 2] - a[0, 2] * a[1, 1]) * d
    t[1, 0] = (a[1, 2] * a[2, 0] - a[1, 0] * a[2, 2]) * d
    t[1, 1] = (a[0, 0] * a[2, 2] - a[0, 2] * a[2, 0]) * d
    t[1, 2] = (a[0, 2] * a[1, 0] - a[0, 0] * a[1, 2]) * d
    t[2, 0] = (a[1, 0] * a[2, 1] - a[1, 1] * a[2, 0]) * d
    t[2, 1] = (a[0, 1] * a[2, 0] - a[0, 0] * a[2, 1]) * d
    t[2, 2] = (a[0, 0] * a[1, 1] - a[0, 1] * a[1, 0]) * d
    return t


def matmul3x3(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """Row-by-column 3x3 product with a fixed left-to-right summation order (so that the
    product library, which does the same on the host, gets the same bits)."""
    c = np.empty((3, 3))
    for i in range(3):
        for j in range(3):
            c[i, j] = a[i, 0] * b[0, j] + a[i, 1] * b[1, j] + a[i, 2] * b[2, j]
    return c


def _cv_round_i32(a: np.ndarray) -> np.ndarray:
    """cvRound on x86: round-half-even; anything that does not fit int32 (or NaN) becomes
    INT_MIN (the 'integer indefinite' value of cvtsd2si).  Verified against cv2 with absurd
    calibrations (oracle/gen_golden.py: 'absurd' cases)."""
    with np.errstate(invalid="ignore"):
        r = np.rint(a)
        bad = ~((r >= -2147483648.0) & (r <= 2147483647.0))
    return np.where(bad, -2147483648.0, r).astype(np.int64)


def init_undistort_rectify_map(K, D, R, P, size):
    """-> (map1 int16 [H,W,2] = (x,y) integer source coords, map2 uint16 [H,W] = (fy<<5)|fx).

    K: 3x3 camera matrix, D: 4/5/8/12 distortion coefficients (k1 k2 p1 p2 k3 k4 k5 k6 s1..s4),
    R: 3x3 rectification rotation, P: 3x4 (or 3x3) new projection, size = (W, H).
    All arithmetic float64, operation order as in OpenCV's per-pixel loop.
    """
    W, H = int(size[0]), int(size[1])
    K = np.asarray(K, np.float64)
    d = np.zeros(14)
    dd = np.asarray(D, np.float64).ravel()
    d[:dd.size] = dd
    k1, k2, p1, p2, k3, k4, k5, k6, s1, s2, s3, s4 = d[:12]
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    iR = invert3x3(matmul3x3(np.asarray(P, np.float64)[:3, :3], np.asarray(R, np.float64)))
    j = np.arange(W, dtype=np.float64)[None, :]
    i = np.arange(H, dtype=np.float64)[:, None]
    with np.errstate(all="ignore"):
        X = iR[0, 0] * j + iR[0, 1] * i + iR[0, 2]
        Y = iR[1, 0] * j + iR[1, 1] * i + iR[1, 2]
        Wd = iR[2, 0] * j + iR[2, 1] * i + iR[2, 2]
        w = 1.0 / Wd
        x = X * w
        y = Y * w
        x2 = x * x
        y2 = y * y
        r2 = x2 + y2
        _2xy = 2 * x * y
        kr = (1 + ((k3 * r2 + k2) * r2 + k1) * r2) / (1 + ((k6 * r2 + k5) * r2 + k4) * r2)
        xd = x * kr + p1 * _2xy + p2 * (r2 + 2 * x2) + s1 * r2 + s2 * r2 * r2
        yd = y * kr + p1 * (r2 + 2 * y2) + p2 * _2xy + s3 * r2 + s4 * r2 * r2
        u = fx * xd + cx
        v = fy * yd + cy
        iu = _cv_round_i32(u * INTER_TAB_SIZE)
        iv = _cv_round_i32(v * INTER_TAB_SIZE)
    sat16 = lambda a: np.clip(a, -32768, 32767).astype(np.int16)
    map1 = np.stack([sat16(iu >> INTER_BITS), sat16(iv >> INTER_BITS)], axis=-1)
    map2 = ((iv & (INTER_TAB_SIZE - 1)) * INTER_TAB_SIZE + (iu & (INTER_TAB_SIZE - 1))).astype(np.uint16)
    return map1, map2


def rectify_maps(calib: dict, size):
    """Both eyes, exactly the reference's two calls (FastACVNet_plus_Algorithm.cpp:139-140)."""
    l = init_undistort_rectify_map(calib["intrinsic_left"], calib["distCoeffs_left"], calib["R_L"], calib["P1"], size)
    r = init_undistort_rectify_map(calib["intrinsic_right"], calib["distCoeffs_right"], calib["R_R"], calib["P2"], size)
    return l, r


# ----------------------------------------------------------------------------------------------
# a3. cv::remap(src, dst, map1, map2, INTER_LINEAR)  (BORDER_CONSTANT, value 0)
#     call sites FastACVNet_plus_Algorithm.cpp:142-143
# ----------------------------------------------------------------------------------------------
def remap_bilinear(src: np.ndarray, map1: np.ndarray, map2: np.ndarray) -> np.ndarray:
    """src uint8 [Hs,Ws,C]; map1 int16 [H,W,2]; map2 uint16 [H,W] -> uint8 [H,W,C]."""
    Hs, Ws = src.shape[:2]
    s = src.reshape(Hs, Ws, -1).astype(np.int64)
    sx = map1[..., 0].astype(np.int64)
    sy = map1[..., 1].astype(np.int64)
    fx = (map2 & 31).astype(np.int64)
    fy = ((map2 >> 5) & 31).astype(np.int64)
    w = [(32 - fx) * (32 - fy) * 32, fx * (32 - fy) * 32, (32 - fx) * fy * 32, fx * fy * 32]

    def tap(yy, xx):
        ok = (yy >= 0) & (yy < Hs) & (xx >= 0) & (xx < Ws)
        v = s[np.clip(yy, 0, Hs - 1), np.clip(xx, 0, Ws - 1)]
        return v * ok[..., None]

    acc = (tap(sy, sx) * w[0][..., None] + tap(sy, sx + 1) * w[1][..., None]
           + tap(sy + 1, sx) * w[2][..., None] + tap(sy + 1, sx + 1) * w[3][..., None])
    out = (acc + (1 << (INTER_REMAP_COEF_BITS - 1))) >> INTER_REMAP_COEF_BITS
    return out.astype(np.uint8).reshape(map2.shape + src.shape[2:])


def rectify_pair(calib: dict, left: np.ndarray, right: np.ndarray):
    """RectifyImage(left, right) of the reference (FastACVNet_plus_Algorithm.cpp:132-145)."""
    size = (left.shape[1], left.shape[0])
    (l1, l2), (r1, r2) = rectify_maps(calib, size)
    return remap_bilinear(left, l1, l2), remap_bilinear(right, r1, r2)


# ----------------------------------------------------------------------------------------------
# a8 (extension). cv::resize(u8, (w,h), INTER_LINEAR), downscale; oracle := OpenCV's 8U path
# ----------------------------------------------------------------------------------------------
def _resize_axis_tab(D: int, S: int):
    f32 = np.float32
    scale = S / D                                    # double
    d = np.arange(D, dtype=np.float64)
    fx = ((d + 0.5) * scale - 0.5).astype(f32)       # OpenCV: fx = (float)((dx+0.5)*scale_x - 0.5)
    sx = np.floor(fx).astype(np.int64)
    fx = (fx - sx.astype(f32)).astype(f32)
    lo = sx < 0
    fx = np.where(lo, f32(0), fx)
    sx = np.where(lo, 0, sx)
    hi = sx >= S - 1
    fx = np.where(hi, f32(0), fx)
    sx = np.where(hi, S - 1, sx)
    c0 = np.rint(((f32(1) - fx) * f32(2048)).astype(f32)).astype(np.int64)
    c1 = np.rint((fx * f32(2048)).astype(f32)).astype(np.int64)
    return sx, np.minimum(sx + 1, S - 1), c0, c1


def resize_linear_u8(src: np.ndarray, size) -> np.ndarray:
    """Bit-exact restatement of OpenCV's 8-bit INTER_LINEAR resize for DOWNSCALES (upscale is
    unpinned; exact 2x downscale is routed to INTER_AREA by OpenCV and
    yields the same values).  src uint8 [H,W,C] -> uint8 [h,w,C]."""
    w, h = int(size[0]), int(size[1])
    H, W = src.shape[:2]
    if (w, h) == (W, H):
        return src.copy()
    if W % w == 0 and H % h == 0 and W // w == 2 and H // h == 2:
        # INTER_LINEAR with exact 2x decimation is executed as INTER_AREA (fast 2x2 box):
        s = src.astype(np.int64)
        return ((s[0::2, 0::2] + s[0::2, 1::2] + s[1::2, 0::2] + s[1::2, 1::2] + 2) >> 2).astype(np.uint8)
    x0, x1, a0, a1 = _resize_axis_tab(w, W)
    y0, y1, b0, b1 = _resize_axis_tab(h, H)
    s = src.reshape(H, W, -1).astype(np.int64)
    T = s[:, x0] * a0[None, :, None] + s[:, x1] * a1[None, :, None]          # [H,w,C] int32 range
    T0, T1 = T[y0], T[y1]
    out = (((b0[:, None, None] * (T0 >> 4)) >> 16) + ((b1[:, None, None] * (T1 >> 4)) >> 16) + 2) >> 2
    return np.clip(out, 0, 255).astype(np.uint8).reshape((h, w) + src.shape[2:])


# ----------------------------------------------------------------------------------------------
# a4. <Model>_preprocess_kernel  (4 variants)
#     RAFTStereo/src/stereo_preprocess.cu:4-39, CREStereo/src/CREStereo_preprocess.cu:4-39,
#     FastACVNet_plus/src/FastACVNet_plus_preprocess.cu:4-39, HitNet/src/HitNet_preprocess.cu:4-53
# ----------------------------------------------------------------------------------------------
_FASTACV_MEAN = (0.485, 0.456, 0.406)     # double literals in the kernel source (:27-29)
_FASTACV_STD = (0.229, 0.224, 0.225)


def norm_lut(model: int) -> np.ndarray:
    """float32 [3,256]: value of output channel ch (R,G,B order) for byte value c, computed with
    the exact type promotions of the reference kernel bodies."""
    c = np.arange(256, dtype=np.float32)
    lut = np.empty((3, 256), np.float32)
    if model in (RAFTSTEREO, CRESTEREO):
        lut[:] = c                                           # float c0 = v[0]; no scaling
    elif model == HITNET:
        lut[:] = (c.astype(np.float64) / 255.0).astype(np.float32)   # (float)v/255.0 (double literal)
    elif model == FASTACVNET_PLUS:
        t = (c / np.float32(255.0)).astype(np.float32)       # c0/255.0f in f32 ...
        for ch in range(3):                                  # ... then double literals promote
            lut[ch] = ((t.astype(np.float64) - _FASTACV_MEAN[ch]) / _FASTACV_STD[ch]).astype(np.float32)
    else:
        raise ValueError(model)
    return lut


def preprocess(model: int, left_u8: np.ndarray, right_u8: np.ndarray, dtype=np.float32):
    """BGR HWC uint8 -> planar RGB tensors.

    RAFT / CRE / FastACV: returns (left [3,H,W], right [3,H,W]).
    HitNet: returns one [6,H,W] tensor (L_r L_g L_b R_r R_g R_b), HitNet_preprocess.cu:38-51.
    dtype float16 (extension a10) := round-to-nearest-even of the float32 value.
    """
    lut = norm_lut(model)

    def one(img):
        rgb = img[..., ::-1]                                    # bgr -> rgb (:21-24)
        return np.stack([lut[ch][rgb[..., ch]] for ch in range(3)], 0).astype(dtype)

    l, r = one(left_u8), one(right_u8)
    if model == HITNET:
        return np.concatenate([l, r], 0)
    return l, r


# ----------------------------------------------------------------------------------------------
# a5. <Model>_reprojectImageTo3D_kernel (4 variants)
#     RAFTStereo/src/stereo_preprocess.cu:41-68 (negate + write back), CREStereo_preprocess.cu:41-63,
#     FastACVNet_plus_preprocess.cu:41-63, HitNet_preprocess.cu:55-75
# ----------------------------------------------------------------------------------------------
def q_to_f32(Q) -> np.ndarray:
    """Calibrationparam_Q[i] = (float)_Q.val[i]  (TRTFastACVNet_plus.cpp:107-113)."""
    return np.asarray(Q, np.float64).reshape(16).astype(np.float32)


def reproject_dense(model: int, disparity: np.ndarray, left_rect_u8: np.ndarray, Q):
    """-> (points float32 [H*W,6] = x y z r g b, disparity_out float32 [H,W]).

    float32 arithmetic, products and sums individually rounded (no FMA contraction), IEEE
    division.  d == 0 gives inf/NaN, d < 0 negative z: nothing is filtered (count == H*W).
    """
    f32 = np.float32
    q = q_to_f32(Q)
    H, W = disparity.shape
    d = disparity.astype(f32)
    if model == RAFTSTEREO:
        d = -d                                                   # temp=-disparity; written back (:54-55)
    col = np.arange(W, dtype=f32)[None, :]
    row = np.arange(H, dtype=f32)[:, None]
    with np.errstate(all="ignore"):
        w = (q[14] * d).astype(f32)
        pts = np.empty((H, W, 6), f32)
        for i in range(3):
            num = ((q[4 * i] * col).astype(f32) + (q[4 * i + 1] * row).astype(f32)).astype(f32)
            num = (num + (q[4 * i + 3] * f32(1))).astype(f32)
            pts[..., i] = (num / w).astype(f32)
    pts[..., 3] = left_rect_u8[..., 2]
    pts[..., 4] = left_rect_u8[..., 1]
    pts[..., 5] = left_rect_u8[..., 0]
    return pts.reshape(H * W, 6), d


def valid_mask(disparity_used: np.ndarray) -> np.ndarray:
    """a9 predicate on the disparity actually used (after RAFT's sign flip)."""
    with np.errstate(invalid="ignore"):
        return np.isfinite(disparity_used) & (disparity_used > 0)


def reproject_compact(model: int, disparity: np.ndarray, left_rect_u8: np.ndarray, Q):
    """a9 (extension): stable row-major filter of the dense output. -> (points [count,6], count, disparity_out)."""
    pts, d = reproject_dense(model, disparity, left_rect_u8, Q)
    keep = valid_mask(d).reshape(-1)
    return pts[keep], int(keep.sum()), d


# ----------------------------------------------------------------------------------------------
# a6. heatmap()  (FastACVNet_plus/test/main.cpp:7-24)
# ----------------------------------------------------------------------------------------------
_JET_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "jet_lut_bgr.txt")


def jet_lut_bgr() -> np.ndarray:
    """uint8 [256,3]: cv::applyColorMap(COLORMAP_JET) table in BGR order; generated once from
    cv2 by oracle/gen_golden.py and committed as oracle/jet_lut_bgr.txt."""
    return np.loadtxt(_JET_PATH, dtype=np.int64).astype(np.uint8).reshape(256, 3)


def heatmap_index(disparity: np.ndarray) -> np.ndarray:
    f32 = np.float32
    d = disparity.astype(f32)
    mn = float(d.min())                                   # minMaxLoc returns doubles
    mx = float(d.max())
    s = f32((mx - mn) / 255)                              # Mat(..., CV_32FC1, (max-min)/255)
    with np.errstate(all="ignore"):
        n = ((d - f32(mn)).astype(f32) / s).astype(f32)   # (disparity - mean_mat) / std_mat
        a = np.rint(np.abs(n))                            # convertScaleAbs: |x| -> cvRound -> saturate
    a = np.where(np.isfinite(a), a, 0.0)                  # NaN (0/0 when max==min) -> INT_MIN -> 0
    return np.clip(a, 0, 255).astype(np.uint8)


def heatmap(disparity: np.ndarray) -> np.ndarray:
    """float32 [H,W] -> uint8 [H,W,3] BGR."""
    return jet_lut_bgr()[heatmap_index(disparity)]


# ----------------------------------------------------------------------------------------------
# f2. stereoRectify: Stereo_Calibration/Stereo_Calibration.cpp:162 calls
#     stereoRectify(K_l, D_l, K_r, D_r, size, R, T, R_L, R_R, P1, P2, Q, CALIB_ZERO_DISPARITY, 0, size, &roiL, &roiR);
#     OpenCV is not vendored in the reference: this restates cv::stereoRectify of OpenCV 4.13 (Bouguet) and is pinned
#     against cv2.stereoRectify by oracle/gen_golden.py -> tests/golden/stereo_rectify.json.
# ----------------------------------------------------------------------------------------------
CALIB_ZERO_DISPARITY = 1024


def rodrigues_to_matrix(r) -> np.ndarray:
    r = np.asarray(r, np.float64).reshape(3)
    theta = math.sqrt(float(r @ r))
    if theta < np.finfo(np.float64).eps:
        return np.eye(3)
    c, s = math.cos(theta), math.sin(theta)
    x, y, z = r / theta
    rrt = np.array([[x * x, x * y, x * z], [x * y, y * y, y * z], [x * z, y * z, z * z]])
    rx = np.array([[0, -z, y], [z, 0, -x], [-y, x, 0]])
    return c * np.eye(3) + (1 - c) * rrt + s * rx


def rodrigues_to_vector(R) -> np.ndarray:
    """cv::Rodrigues matrix -> vector: nearest rotation (U Vt) first, then axis * angle (general case only)."""
    u, _, vt = np.linalg.svd(np.asarray(R, np.float64).reshape(3, 3))
    R = u @ vt
    r = np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]])
    s = math.sqrt(float(r @ r) * 0.25)
    c = min(1.0, max(-1.0, (R[0, 0] + R[1, 1] + R[2, 2] - 1) * 0.5))
    theta = math.acos(c)
    if s < 1e-5:
        if c > 0:
            return np.zeros(3)
        raise ValueError("rotation by pi: not needed for a stereo rig")
    return r * (1.0 / (2 * s) * theta)


def _undistort_point(u, v, K, k, RR=None):
    """cv::undistortPoints for one point, default criteria = 5 iterations; k = 14 coefficients."""
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    ifx, ify = 1.0 / fx, 1.0 / fy
    x = (u - cx) * ifx
    y = (v - cy) * ify
    x0, y0 = x, y
    for _ in range(5):
        r2 = x * x + y * y
        icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2)
        if icdist < 0:
            x, y = (u - cx) * ifx, (v - cy) * ify
            break
        dx = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2
        dy = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2
        x = (x0 - dx) * icdist
        y = (y0 - dy) * icdist
    if RR is not None:
        xx = RR[0, 0] * x + RR[0, 1] * y + RR[0, 2]
        yy = RR[1, 0] * x + RR[1, 1] * y + RR[1, 2]
        ww = 1.0 / (RR[2, 0] * x + RR[2, 1] * y + RR[2, 2])
        x, y = xx * ww, yy * ww
    return x, y


def _mat3(a, b):
    """3x3 product with OpenCV's summation order (no BLAS, no FMA)."""
    return np.array([[sum(float(a[i, k]) * float(b[k, j]) for k in range(3)) for j in range(3)] for i in range(3)])


def _undistort_rectangles(K, k, R, P, size):
    N = 9
    w, h = size
    RR = _mat3(np.asarray(P, np.float64)[:3, :3], R)
    FM = float(np.finfo(np.float32).max)
    iX0, iX1, iY0, iY1 = -FM, FM, -FM, FM
    oX0, oX1, oY0, oY1 = FM, -FM, FM, -FM
    for y in range(N):
        for x in range(N):
            px, py = _undistort_point(float(x) * (w - 1) / (N - 1), float(y) * (h - 1) / (N - 1), K, k, RR)
            oX0, oX1, oY0, oY1 = min(oX0, px), max(oX1, px), min(oY0, py), max(oY1, py)
            if x == 0:
                iX0 = max(iX0, px)
            if x == N - 1:
                iX1 = min(iX1, px)
            if y == 0:
                iY0 = max(iY0, py)
            if y == N - 1:
                iY1 = min(iY1, py)
    return (iX0, iY0, iX1 - iX0, iY1 - iY0), (oX0, oY0, oX1 - oX0, oY1 - oY0)


def stereo_rectify(K1, D1, K2, D2, size, R, T, flags=CALIB_ZERO_DISPARITY, alpha=-1.0, new_size=(0, 0)):
    """-> R1, R2 (3x3), P1, P2 (3x4), Q (4x4), roi1, roi2 (x, y, w, h)."""
    f32 = np.float32
    K1, K2 = np.asarray(K1, np.float64).reshape(3, 3), np.asarray(K2, np.float64).reshape(3, 3)
    ks = []
    for D in (D1, D2):
        k = np.zeros(14)
        d = np.asarray(D, np.float64).ravel()
        k[:len(d)] = d
        ks.append(k)
    nx, ny = size
    T = np.asarray(T, np.float64).reshape(3)
    r_r = rodrigues_to_matrix(rodrigues_to_vector(R) * -0.5)          # each camera turns half way
    t = np.array([sum(float(r_r[i, j]) * float(T[j]) for j in range(3)) for i in range(3)])
    idx = 0 if abs(t[0]) > abs(t[1]) else 1
    c, nt = t[idx], math.sqrt(float(t[0] * t[0] + t[1] * t[1] + t[2] * t[2]))
    uu = np.zeros(3)
    uu[idx] = 1 if c > 0 else -1
    ww = np.array([t[1] * uu[2] - t[2] * uu[1], t[2] * uu[0] - t[0] * uu[2], t[0] * uu[1] - t[1] * uu[0]])
    nw = math.sqrt(float(ww[0] * ww[0] + ww[1] * ww[1] + ww[2] * ww[2]))
    if nw > 0:
        ww = ww * (math.acos(abs(c) / nt) / nw)
    wR = rodrigues_to_matrix(ww)
    R1, R2 = _mat3(wR, r_r.T), _mat3(wR, r_r)
    t = np.array([sum(float(R2[i, j]) * float(T[j]) for j in range(3)) for i in range(3)])
    W, H = new_size if new_size[0] * new_size[1] != 0 else size
    ratio = (W / nx / 2) if idx == 1 else (H / ny / 2)
    fc = (K1[idx ^ 1, idx ^ 1] + K2[idx ^ 1, idx ^ 1]) * ratio
    cc = []
    for K, k, Rk in ((K1, ks[0], R1), (K2, ks[1], R2)):
        ax = ay = 0.0
        for i in range(4):                                            # image corners; OpenCV keeps them as float
            x, y = _undistort_point(float(f32((i % 2) * (nx - 1))), float(f32((0 if i < 2 else 1) * (ny - 1))), K, k)
            X = (float(f32(x)), float(f32(y)), 1.0)
            Y = [Rk[j, 0] * X[0] + Rk[j, 1] * X[1] + Rk[j, 2] * X[2] for j in range(3)]
            z = 1.0 / Y[2] if Y[2] != 0 else 1.0
            ax += float(f32(fc * (Y[0] * z)))
            ay += float(f32(fc * (Y[1] * z)))
        cc.append([(nx - 1) / 2 - ax / 4, (ny - 1) / 2 - ay / 4])
    if flags & CALIB_ZERO_DISPARITY:
        cc[0][0] = cc[1][0] = (cc[0][0] + cc[1][0]) * 0.5
        cc[0][1] = cc[1][1] = (cc[0][1] + cc[1][1]) * 0.5
    elif idx == 0:
        cc[0][1] = cc[1][1] = (cc[0][1] + cc[1][1]) * 0.5
    else:
        cc[0][0] = cc[1][0] = (cc[0][0] + cc[1][0]) * 0.5
    P1 = np.zeros((3, 4))
    P1[0, 0] = P1[1, 1] = fc
    P1[0, 2], P1[1, 2], P1[2, 2] = cc[0][0], cc[0][1], 1
    P2 = P1.copy()
    P2[0, 2], P2[1, 2] = cc[1][0], cc[1][1]
    P2[idx, 3] = t[idx] * fc
    alpha = min(alpha, 1.0)
    in1, out1 = _undistort_rectangles(K1, ks[0], R1, P1, size)
    in2, out2 = _undistort_rectangles(K2, ks[1], R2, P2, size)
    (cx1_0, cy1_0), (cx2_0, cy2_0) = cc
    cx1, cy1, cx2, cy2 = W * cx1_0 / nx, H * cy1_0 / ny, W * cx2_0 / nx, H * cy2_0 / ny
    s = 1.0
    if alpha >= 0:
        def ext(cx, cy, cx0, cy0, r, fn):
            x, y, w, h = r
            return fn(fn(fn(cx / (cx0 - x), cy / (cy0 - y)), (W - 1 - cx) / (x + w - cx0)), (H - 1 - cy) / (y + h - cy0))
        s0 = max(ext(cx2, cy2, cx2_0, cy2_0, in2, max), ext(cx1, cy1, cx1_0, cy1_0, in1, max))
        s1 = min(ext(cx2, cy2, cx2_0, cy2_0, out2, min), ext(cx1, cy1, cx1_0, cy1_0, out1, min))
        s = s0 * (1 - alpha) + s1 * alpha
    fc *= s
    P1[0, 0] = P1[1, 1] = P2[0, 0] = P2[1, 1] = fc
    P1[0, 2], P1[1, 2], P2[0, 2], P2[1, 2] = cx1, cy1, cx2, cy2
    P2[idx, 3] = s * P2[idx, 3]

    def roi(inn, cx0, cy0, cx, cy):
        x, y, w, h = inn
        rx, ry = math.ceil((x - cx0) * s + cx), math.ceil((y - cy0) * s + cy)
        rw, rh = math.floor(w * s), math.floor(h * s)
        x0, y0, x1, y1 = max(rx, 0), max(ry, 0), min(rx + rw, W), min(ry + rh, H)
        return (0, 0, 0, 0) if x1 <= x0 or y1 <= y0 else (x0, y0, x1 - x0, y1 - y0)
    Q = np.array([[1, 0, 0, -cx1], [0, 1, 0, -cy1], [0, 0, 0, fc], [0, 0, -1.0 / t[idx], ((cx1 - cx2) if idx == 0 else (cy1 - cy2)) / t[idx]]])
    return R1, R2, P1, P2, Q, roi(in1, cx1_0, cy1_0, cx1, cy1), roi(in2, cx2_0, cy2_0, cx2, cy2)



def disparity_u8(disparity: np.ndarray) -> np.ndarray:
    """What cv::imwrite("disparity.jpg", disparity) encodes (FastACVNet_plus/test/main.cpp:55): imwrite converts a
    CV_32FC1 Mat with convertTo(CV_8U) = saturate_cast<uchar>(cvRound(d)); cvRound (cvtss2si) rounds half to even and
    returns INT_MIN for NaN and |d| >= 2^31, which saturates to 0.  Pinned against cv2 (PNG round trip) in gen_golden.py."""
    d = np.asarray(disparity, np.float32)
    with np.errstate(all="ignore"):
        r = np.rint(d.astype(np.float64))
        ok = np.abs(d) < np.float32(2147483648.0)          # False for NaN
    return np.where(ok, np.clip(np.where(ok, r, 0.0), 0, 255), 0).astype(np.uint8)


# ----------------------------------------------------------------------------------------------
# synthetic inputs - deterministic, shared by tests and bench baselines
# ----------------------------------------------------------------------------------------------
def synthetic_disparity(W: int, H: int, seed: int = 1234, invalid: bool = True) -> np.ndarray:
    c = np.arange(W, dtype=np.float64)[None, :]
    r = np.arange(H, dtype=np.float64)[:, None]
    d = 4 + 60 * (0.5 + 0.5 * np.sin(2 * math.pi * 3 * c / W)) * (0.5 + 0.5 * np.cos(2 * math.pi * 2 * r / H))
    d = d.astype(np.float32)
    if invalid:
        rng = np.random.default_rng(seed)
        u = rng.random((H, W))
        d[u < 0.05] = 0.0
        d[(u >= 0.05) & (u < 0.06)] = -1.0
    return d


def synthetic_image(W: int, H: int, seed: int) -> np.ndarray:
    return np.random.default_rng(seed).integers(0, 256, size=(H, W, 3), dtype=np.uint8)


def pointcloud_txt_lines(points: np.ndarray, limit: int | None = None):
    """pointcloud.txt of the demo (FastACVNet_plus/test/main.cpp:62-69): default
    ``operator<<(float)`` formatting == printf("%g") with 6 significant digits."""
    pts = points if limit is None else points[:limit]
    return [" ".join(("%g" % v) for v in p) for p in pts]


# ----------------------------------------------------------------------------------------------
# whole path, one pair  (RunFastACVNet_plus_RectifyImage, FastACVNet_plus_Algorithm.cpp:59-75)
# ----------------------------------------------------------------------------------------------
def run_pair(model: int, calib: dict, left: np.ndarray, right: np.ndarray, disparity_from_net: np.ndarray,
             rectify: bool = True, net_size=None, dtype=np.float32):
    """Returns dict(rect_left, rect_right, tensors, points, disparity, heatmap)."""
    if rectify:
        rl, rr = rectify_pair(calib, left, right)
    else:
        rl, rr = left, right
    nl, nr = rl, rr
    Q = np.asarray(calib["Q"], np.float64)
    if net_size is not None and tuple(net_size) != (left.shape[1], left.shape[0]):
        nl, nr = resize_linear_u8(rl, net_size), resize_linear_u8(rr, net_size)
    tensors = preprocess(model, nl, nr, dtype)
    pts, d = reproject_dense(model, disparity_from_net, nl, Q)
    return dict(rect_left=rl, rect_right=rr, net_left=nl, net_right=nr, tensors=tensors, points=pts,
                disparity=d, heatmap=heatmap(d))
