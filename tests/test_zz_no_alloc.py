"""sa_reproject_compact_noalloc against sa_reproject(SA_COMPACT).  Needs a B200: -m gpu.

The no-alloc entry point exists for callers that cannot allocate device memory where they need the point cloud (a
TensorRT plugin's enqueue(), a real-time callback that must not allocate after start-up, a process whose own pool owns
the GPU): its compaction scratch is an array built into the library instead of buffers a handle allocated.  It runs the
same kernels, so its output must equal the handle-based entry point's - checked here with the tolerance the other
reprojection tests use (1e-5 relative on XYZ, counts / order / colours exact), on two calls in a row with different
data so that whatever the first call left in the shared scratch cannot leak into the second.  So that it does not stand
on another CUDA path alone, every result is also compared with the CPU oracle's compaction (oracle.reproject_compact),
and a second test drives the options the first leaves at their defaults: a caller-owned stream, a left image with padded
rows, explicit frame strides.

This file is named to be collected last: if it fails under `pytest -x`, every other test has already run.
"""
import ctypes as C

import numpy as np
import pytest

from oracle import oracle as O

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
                # and against the CPU oracle, so the entry point is checked by something that is not a CUDA kernel
                o_pts, o_cnt, o_d = O.reproject_compact(O.RAFTSTEREO if negate else O.FASTACVNET_PLUS, d[i], img[i], Q)
                assert o_cnt == k
                _close(got_pts[i, :k], o_pts)
                assert np.array_equal(got_pts[i, :k, 3:], o_pts[:, 3:])
                assert np.array_equal(got_d[i], o_d, equal_nan=True)


def test_noalloc_caller_stream_pitched_left_image_and_frame_strides():
    """The arguments the first test leaves at 0 / NULL: work queued on the caller's own stream (nothing synchronised by the
    library), left images whose rows are padded (a region of a wider capture buffer), disparity frames further apart than one
    frame (plane 0 of a two-plane network output).  640x480 takes the vector paths.  Checked against the CPU oracle, twice."""
    import torch
    from stereoalgorithms_b200 import _lib as L
    lib = L.load()
    torch.cuda.set_device(0)
    n, W, H = 4, 640, 480
    pitch = 3 * W + 64
    Q = np.array([[1, 0, 0, -310.5], [0, 1, 0, -251.25], [0, 0, 0, 517.8], [0, 0, 0.0166, 0]], np.float64)
    qf = np.ascontiguousarray(Q.reshape(16).astype(np.float32))
    stream = torch.cuda.Stream()
    for call in range(2):
        img, d = _inputs(300 + call, n, W, H)
        wide = np.full((n, H, pitch), 0xAB, np.uint8)                     # padding bytes that must never show up as colours
        wide[:, :, :3 * W] = img.reshape(n, H, 3 * W)
        two_plane = np.full((n, 2, H, W), np.nan, np.float32)
        two_plane[:, 0] = d
        with torch.cuda.stream(stream):                                   # inputs are produced on the caller's stream too
            dd = torch.from_numpy(two_plane).cuda(non_blocking=False)
            di = torch.from_numpy(wide).cuda()
            pts = torch.zeros((n, W * H, 6), dtype=torch.float32, device="cuda")
            cnt = torch.full((n,), -1, dtype=torch.int32, device="cuda")
            err = C.create_string_buffer(256)
            rc = lib.sa_reproject_compact_noalloc(dd.data_ptr(), 2 * W * H, di.data_ptr(), pitch, pitch * H, qf.ctypes.data, W, H, n, 0,
                                                  pts.data_ptr(), cnt.data_ptr(), stream.cuda_stream, err, 256)
            assert rc == 0, err.value
        stream.synchronize()
        got_cnt, got_pts, got_d = cnt.cpu().numpy(), pts.cpu().numpy(), dd.cpu().numpy()
        assert np.array_equal(got_d[:, 0], d, equal_nan=True) and np.isnan(got_d[:, 1]).all()   # plane 1 untouched
        for i in range(n):
            o_pts, o_cnt, _ = O.reproject_compact(O.FASTACVNET_PLUS, d[i], img[i], Q)
            assert got_cnt[i] == o_cnt
            _close(got_pts[i, :o_cnt], o_pts)
            assert np.array_equal(got_pts[i, :o_cnt, 3:], o_pts[:, 3:])
            assert not got_pts[i, o_cnt:].any()


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
    # tiny frames pass the slot limit with more frames than a grid has rows: still a clear SA_ERR_TOO_LARGE, not a launch error
    rc = lib.sa_reproject_compact_noalloc(buf.data_ptr(), 0, buf.data_ptr(), 0, 0, q.ctypes.data, 32, 32, 65536, 0,
                                          buf.data_ptr(), buf.data_ptr(), None, err, 512)
    assert rc == L.SA_ERR_TOO_LARGE and b"65535" in err.value
    # the point rows are written as float2: a destination at a 4-byte offset is refused instead of faulting
    rc = lib.sa_reproject_compact_noalloc(buf.data_ptr(), 0, buf.data_ptr(), 0, 0, q.ctypes.data, 8, 8, 1, 0,
                                          buf.data_ptr() + 4, buf.data_ptr(), None, err, 512)
    assert rc == L.SA_ERR_INVALID_ARG and b"aligned" in err.value
    rc = lib.sa_reproject_compact_noalloc(None, 0, buf.data_ptr(), 0, 0, q.ctypes.data, 64, 16, 1, 0, buf.data_ptr(), buf.data_ptr(), None, err, 512)
    assert rc == L.SA_ERR_INVALID_ARG and b"null" in err.value
    rc = lib.sa_reproject_compact_noalloc(buf.data_ptr(), 0, buf.data_ptr(), 10, 0, q.ctypes.data, 64, 16, 1, 0, buf.data_ptr(), buf.data_ptr(), None, err, 512)
    assert rc == L.SA_ERR_INVALID_ARG and b"left_pitch" in err.value
