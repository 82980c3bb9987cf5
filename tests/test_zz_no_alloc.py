"""sa_reproject_compact_noalloc against sa_reproject(SA_COMPACT).  Needs a B200: -m gpu.

The no-alloc entry point exists for callers that cannot allocate device memory where they need the point cloud (a
TensorRT plugin's enqueue(), a real-time callback that must not allocate after start-up, a process whose own pool owns
the GPU): its compaction scratch is an array built into the library instead of buffers a handle allocated.  It runs the
same kernels, so its output must equal the handle-based entry point's - checked here with the tolerance the other
reprojection tests use (1e-5 relative on XYZ, counts / order / colours exact), on two calls in a row with different
data so that whatever the first call left in the shared scratch cannot leak into the second.

This file is named to be collected last: if it fails under `pytest -x`, every other test has already run.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-5


def _inputs(seed, n, W, H):
    rng = np.random.default_rng(seed)
    img = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    d = rng.normal(25.0, 20.0, (n, H, W)).astype(np.float32)          # plenty of negatives
    d[rng.random(d.shape) < 0.07] = 0.0                                # and zeros
    d[rng.random(d.shape) < 0.001] = np.inf
    d[rng.random(d.shape) < 0.001] = np.nan
    return img, d


def _close(a, b):
    assert np.array_equal(np.isnan(a), np.isnan(b)) and np.array_equal(np.isposinf(a), np.isposinf(b)) and np.array_equal(np.isneginf(a), np.isneginf(b))
    f = np.isfinite(b)
    assert np.all(np.abs(a[f] - b[f]) <= RTOL * np.maximum(np.abs(a[f]), np.abs(b[f])))


@pytest.mark.parametrize("negate", [0, 1])
def test_noalloc_matches_handle_entry_point_twice(negate):
    import torch
    from stereoalgorithms_b200 import _lib as L
    from stereoalgorithms_b200.geometry import COMPACT, StereoGeometry
    lib = L.load()
    n, W, H = 3, 322, 241                         # odd size: scalar paths, a partial last CTA per frame
    Q = np.array([[1, 0, 0, -160.5], [0, 1, 0, -120.25], [0, 0, 0, 270.8], [0, 0, 0.0166, 0]], np.float64)
    qf = np.ascontiguousarray(Q.reshape(16).astype(np.float32))
    model = "RAFTStereo" if negate else "FastACVNet_plus"
    torch.cuda.set_device(0)
    with StereoGeometry(W, H, model=model, max_batch=n) as g:
        g.set_Q(Q)
        for call in range(2):                     # two calls in a row, different data
            img, d = _inputs(100 + call, n, W, H)
            # existing entry point
            g.rectify_preprocess(img, img, rectify=False)
            g.set_disparity(d[:, None])
            want_pts, want_cnt = g.reproject(n, COMPACT)
            want_d = g.disparity(n)
            # new entry point: caller-owned device buffers only
            dd = torch.from_numpy(d.copy()).cuda()
            di = torch.from_numpy(img).cuda()
            pts = torch.zeros((n, W * H, 6), dtype=torch.float32, device="cuda")
            cnt = torch.full((n,), -1, dtype=torch.int32, device="cuda")
            err = C.create_string_buffer(256)
            torch.cuda.synchronize()
            rc = lib.sa_reproject_compact_noalloc(dd.data_ptr(), 0, di.data_ptr(), 0, 0, qf.ctypes.data, W, H, n, negate,
                                                  pts.data_ptr(), cnt.data_ptr(), None, err, 256)
            assert rc == 0, err.value
            torch.cuda.synchronize()
            got_cnt, got_pts, got_d = cnt.cpu().numpy(), pts.cpu().numpy(), dd.cpu().numpy()
            assert np.array_equal(got_cnt, want_cnt)
            assert np.array_equal(got_d, want_d, equal_nan=True)          # RAFT: flipped in place, like the handle path
            for i in range(n):
                k = int(want_cnt[i])
                assert 0 < k < W * H
                _close(got_pts[i, :k], want_pts[i, :k])
                assert np.array_equal(got_pts[i, :k, 3:], want_pts[i, :k, 3:])   # same pixels, same order
                assert not got_pts[i, k:].any()


def test_noalloc_rejects_oversize_and_bad_arguments():
    import torch
    from stereoalgorithms_b200 import _lib as L
    lib = L.load()
    torch.cuda.set_device(0)
    buf = torch.zeros(1024, dtype=torch.float32, device="cuda")
    q = np.eye(4, dtype=np.float32).reshape(16)
    err = C.create_string_buffer(512)
    # 3840x2160 needs 8100 slots per frame: 9 frames exceed the 65536-slot built-in scratch.  Rejected before any launch,
    # so the (too small) dummy buffers are never touched.
    rc = lib.sa_reproject_compact_noalloc(buf.data_ptr(), 0, buf.data_ptr(), 0, 0, q.ctypes.data, 3840, 2160, 9, 0,
                                          buf.data_ptr(), buf.data_ptr(), None, err, 512)
    assert rc == L.SA_ERR_TOO_LARGE
    assert b"65536" in err.value and b"72900" in err.value
    rc = lib.sa_reproject_compact_noalloc(None, 0, buf.data_ptr(), 0, 0, q.ctypes.data, 64, 16, 1, 0, buf.data_ptr(), buf.data_ptr(), None, err, 512)
    assert rc == L.SA_ERR_INVALID_ARG and b"null" in err.value
    rc = lib.sa_reproject_compact_noalloc(buf.data_ptr(), 0, buf.data_ptr(), 10, 0, q.ctypes.data, 64, 16, 1, 0, buf.data_ptr(), buf.data_ptr(), None, err, 512)
    assert rc == L.SA_ERR_INVALID_ARG and b"left_pitch" in err.value
