"""The reference-named entry points (Initialize / Run<Model>[_RectifyImage] / Version / Release), used the way the
reference's demo programs use them (FastACVNet_plus/test/main.cpp:27-74).  Needs a B200: -m gpu."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from oracle import oracle as O

pytestmark = pytest.mark.gpu

CALIB = os.path.join(GOLDEN, "calib_b.yml")     # the */test/StereoCalibration.yml the demos load


def _check_points(a, b):
    assert np.array_equal(np.isnan(a), np.isnan(b)) and np.array_equal(np.isinf(a), np.isinf(b))
    f = np.isfinite(b)
    assert np.all(np.abs(a[f] - b[f]) <= 1e-5 * np.maximum(np.abs(a[f]), np.abs(b[f])))


def test_demo_main_fastacvnet_plus(demo_pair, calibs, tmp_path):
    from stereoalgorithms_b200 import compat
    from stereoalgorithms_b200.compat import FastACVNet_plus as F, Mat
    imageL, imageR = demo_pair
    fastacvnet = F.Initialize("fast_acvnet_plus_generalization_opset16_480x640.onnx", 0, CALIB)
    assert fastacvnet
    assert F.Version(fastacvnet) == "FastACVNet_plus_Algorithm_V1.0"
    pointcloud = np.empty(imageL.shape[1] * imageL.shape[0] * 6, np.float32)
    disparity = Mat()
    for _ in range(5):
        imageL1, imageR1 = imageL.copy(), imageR.copy()
        assert F.RunFastACVNet_plus_RectifyImage(fastacvnet, imageL1, imageR1, pointcloud, disparity) == 0
    # left/right now hold the rectified images, as after cv::remap(img, img, ...)
    c = calibs["calib_b"]
    rl, rr = O.rectify_pair(c, imageL, imageR)
    assert np.array_equal(imageL1, rl) and np.array_equal(imageR1, rr)
    assert disparity.array.shape == (480, 640) and disparity.array.dtype == np.float32
    want, _ = O.reproject_dense(O.FASTACVNET_PLUS, disparity.array, rl, c["Q"])
    _check_points(pointcloud.reshape(-1, 6), want)
    hm = compat.heatmap(fastacvnet, disparity)
    assert np.array_equal(hm, O.heatmap(disparity.array))
    # pointcloud.txt, first lines
    path = str(tmp_path / "pointcloud.txt")
    compat.write_pointcloud_txt(path, pointcloud[:6 * 1000])
    with open(path) as f:
        lines = f.read().splitlines()
    assert lines == O.pointcloud_txt_lines(pointcloud.reshape(-1, 6), 1000)
    # the no-rectify flavour leaves the images alone
    a, b = imageL.copy(), imageR.copy()
    assert F.RunFastACVNet_plus(fastacvnet, a, b, pointcloud, disparity) == 0
    assert np.array_equal(a, imageL) and np.array_equal(b, imageR)
    want, _ = O.reproject_dense(O.FASTACVNET_PLUS, disparity.array, imageL, c["Q"])
    _check_points(pointcloud.reshape(-1, 6), want)
    assert F.Release(fastacvnet) == 0


def test_reused_pointcloud_buffer_is_page_locked_and_results_do_not_change(demo_pair, calibs):
    """The demos pass the same `pointcloud` to every Run (FastACVNet_plus/test/main.cpp:40-50); with SetHostRegister(1) the
    library page-locks it from the second call on.  Results must not depend on that, on switching buffers, or on switching the feature off."""
    from stereoalgorithms_b200 import _lib as L
    from stereoalgorithms_b200.compat import FastACVNet_plus as F, Mat
    lib = L.load()
    imageL, imageR = demo_pair
    p = F.Initialize("m.onnx", 0, CALIB)
    assert F.SetHostRegister(p, True) == 0                   # opt-in: the caller promises not to free a pinned buffer early
    n = imageL.shape[0] * imageL.shape[1] * 6
    bufs = [np.zeros(n, np.float32), np.zeros(n, np.float32)]
    d = Mat()
    outs = []
    for buf_i in (0, 0, 0, 1, 1, 0, 0):                      # pinned from call 2; dropped and re-pinned as the buffer changes
        bufs[buf_i][:] = 0
        assert F.RunFastACVNet_plus_RectifyImage(p, imageL.copy(), imageR.copy(), bufs[buf_i], d) == 0
        outs.append(bufs[buf_i].copy())
    for o in outs[1:]:
        assert np.array_equal(o, outs[0], equal_nan=True)
    # a buffer the caller page-locked already: registration fails inside, the call still works
    own = np.zeros(n, np.float32)
    assert lib.sa_host_register(own.ctypes.data, own.nbytes) == 0
    for _ in range(3):
        assert F.RunFastACVNet_plus_RectifyImage(p, imageL.copy(), imageR.copy(), own, d) == 0
    assert np.array_equal(own, outs[0], equal_nan=True)
    assert lib.sa_host_unregister(own.ctypes.data) == 0
    assert F.SetHostRegister(p, False) == 0
    for _ in range(3):
        bufs[0][:] = 0
        assert F.RunFastACVNet_plus_RectifyImage(p, imageL.copy(), imageR.copy(), bufs[0], d) == 0
    assert np.array_equal(bufs[0], outs[0], equal_nan=True)
    assert F.Release(p) == 0
    # released: the buffer is ordinary memory again and can be registered by someone else
    assert lib.sa_host_register(bufs[0].ctypes.data, bufs[0].nbytes) == 0 and lib.sa_host_unregister(bufs[0].ctypes.data) == 0


@pytest.mark.parametrize("name", ["RAFTStereo", "HitNet", "CREStereo"])
def test_other_models(demo_pair, calibs, name):
    from stereoalgorithms_b200 import compat
    M = getattr(compat, name)
    run = {"RAFTStereo": "RunRAFTStereo", "HitNet": "RunHitNet", "CREStereo": "RunCREStereo_RectifyImage"}[name]
    assert hasattr(M, "RunCREStereo") == (name == "CREStereo") and not hasattr(M, "RunRAFTStereo_RectifyImage")
    p = M.Initialize("model.onnx", 0, CALIB)
    assert M.Version(p) == {"RAFTStereo": "RAFTStereoAlgorithm_V1.0", "HitNet": "HitNetAlgorithm_V1.0", "CREStereo": "CREStereoAlgorithm_V1.0"}[name]
    l, r = demo_pair[0].copy(), demo_pair[1].copy()
    pc = np.empty(640 * 480 * 6, np.float32)
    d = compat.Mat()
    assert getattr(M, run)(p, l, r, pc, d) == 0
    c = calibs["calib_b"]
    rl, rr = O.rectify_pair(c, *demo_pair)
    assert np.array_equal(l, rl) and np.array_equal(r, rr)                 # always-rectify entry points mutate in place
    assert (d.array > 0).mean() > 0.9                                      # RAFT: returned disparity is the flipped, positive one
    want, _ = O.reproject_dense(O.FASTACVNET_PLUS, d.array, rl, c["Q"])    # d is already the disparity as used
    _check_points(pc.reshape(-1, 6), want)
    assert M.Release(p) == 0


def test_error_semantics(demo_pair, tmp_path):
    from stereoalgorithms_b200.compat import CREStereo as M, Mat
    pc = np.empty(640 * 480 * 6, np.float32)
    d = Mat()
    # missing calibration: Initialize still returns a handle (reference: prints a warning and goes on)
    p = M.Initialize("m.onnx", 0, str(tmp_path / "nope.yml"))
    assert p and "calibration_path does not exist" in M.LastError(p)
    l, r = demo_pair[0].copy(), demo_pair[1].copy()
    assert M.RunCREStereo_RectifyImage(p, l, r, pc, d) == -1               # cannot rectify without a calibration
    assert M.RunCREStereo(p, None, r, pc, d) == -1                         # "Image_src is empty!!!"
    assert M.RunCREStereo(p, np.empty((0, 0, 3), np.uint8), r, pc, d) == -1
    assert M.Release(p) == 0
    # failed init (no such GPU) latches: every Run returns -1
    p = M.Initialize("m.onnx", 99, CALIB)
    assert p
    assert M.RunCREStereo(p, l, r, pc, d) == -1
    assert M.Release(p) == 0


def test_other_image_size_than_640x480(calibs):
    """The reference asserts 640x480; here the geometry is re-created for the size it is given."""
    from stereoalgorithms_b200.compat import FastACVNet_plus as F, Mat
    W, H = 320, 240
    img = O.synthetic_image(W, H, 1)
    p = F.Initialize("m.onnx", 0, CALIB)
    pc = np.empty(W * H * 6, np.float32)
    d = Mat()
    assert F.RunFastACVNet_plus(p, img.copy(), img.copy(), pc, d) == 0
    assert d.array.shape == (H, W)
    want, _ = O.reproject_dense(O.FASTACVNET_PLUS, d.array, img, calibs["calib_b"]["Q"])
    _check_points(pc.reshape(-1, 6), want)
    F.Release(p)


def test_cpp_shim_demo_program(demo_pair, calibs, tmp_path):
    """shims/FastACVNet_plus_Algorithm.cpp compiled against the stand-in cv::Mat, driven by a cut-down copy of the
    reference's demo main (tests/shim_demo.cpp): the C++ route into the library, end to end."""
    import subprocess
    from conftest import ROOT
    from test_shims_cpu import build_shim
    so = build_shim("FastACVNet_plus_Algorithm", tmp_path)
    exe = str(tmp_path / "shim_demo")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-I", os.path.join(ROOT, "tests", "fake_opencv"), "-o", exe,
                           os.path.join(ROOT, "tests", "shim_demo.cpp"), so, "-Wl,-rpath," + str(tmp_path)])
    l, r = demo_pair
    l.tofile(str(tmp_path / "l.bgr"))
    r.tofile(str(tmp_path / "r.bgr"))
    out = subprocess.run([exe, CALIB, str(tmp_path / "l.bgr"), str(tmp_path / "r.bgr"), "640", "480", str(tmp_path / "o")],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    assert out.stdout.strip() == "FastACVNet_plus_Algorithm_V1.0"
    c = calibs["calib_b"]
    rl, rr = O.rectify_pair(c, l, r)
    assert np.array_equal(np.fromfile(str(tmp_path / "o_rect_left.bgr"), np.uint8).reshape(480, 640, 3), rl)
    assert np.array_equal(np.fromfile(str(tmp_path / "o_rect_right.bgr"), np.uint8).reshape(480, 640, 3), rr)
    d = np.fromfile(str(tmp_path / "o_disparity.f32"), np.float32).reshape(480, 640)
    pts = np.fromfile(str(tmp_path / "o_pointcloud.f32"), np.float32).reshape(-1, 6)
    want, _ = O.reproject_dense(O.FASTACVNET_PLUS, d, rl, c["Q"])
    _check_points(pts, want)


def test_engine_on_the_reference_named_path_is_ordered_on_the_handle_stream(demo_pair, calibs):
    """sa_ref_SetEngine: the slot of context->enqueueV3(stream) between preprocess and reproject
    (FastACVNet_plus/src/TRTFastACVNet_plus.cpp:133-160), through Run<Model>_RectifyImage.  The callback is handed the
    stream both halves run on.  It behaves like an asynchronous engine: it snapshots the tensors it was given with a
    stream-ordered copy, keeps the stream busy with a long spin kernel, and only then writes the 'network output' - all
    without synchronising.  Stage B must wait for that write (the cloud equals the oracle's for THIS disparity, twice in a
    row with different outputs), and the snapshot must already hold the finished tensors.  With a NULL stream handed to the
    callback - work on the legacy default stream, which the handle's non-blocking stream does not order against - this fails."""
    import torch
    from cuda.bindings import runtime as rt
    from stereoalgorithms_b200.compat import FastACVNet_plus as F, Mat
    c = calibs["calib_b"]
    W, H = 640, 480
    imageL, imageR = demo_pair
    nbytes = W * H * 4
    snap = [torch.zeros(3 * H * W, dtype=torch.float32, device="cuda") for _ in range(2)]
    net_out = torch.zeros(H * W, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    seen = []
    D2D = rt.cudaMemcpyKind.cudaMemcpyDeviceToDevice

    def engine(stream, first_frame, n, left_tensor, right_tensor, disparity):
        s = int(stream or 0)
        seen.append(s)
        ok = rt.cudaError_t.cudaSuccess
        for buf, src in zip(snap, (left_tensor, right_tensor)):
            if rt.cudaMemcpyAsync(buf.data_ptr(), src, 3 * nbytes, D2D, s)[0] != ok:
                return 1
        with torch.cuda.stream(torch.cuda.ExternalStream(s)):
            torch.cuda._sleep(100_000_000)                               # ~50 ms of spinning on the stream we were given
        return 0 if rt.cudaMemcpyAsync(disparity, net_out.data_ptr(), nbytes, D2D, s)[0] == ok else 1

    p = F.Initialize("m.onnx", 0, CALIB)
    assert F.SetEngine(p, engine) == 0
    rl, rr = O.rectify_pair(c, imageL, imageR)
    tl, tr = O.preprocess(O.FASTACVNET_PLUS, rl, rr)
    pointcloud = np.empty(W * H * 6, np.float32)
    disparity = Mat()
    for k in range(2):
        D = O.synthetic_disparity(W, H, 500 + k)
        net_out.copy_(torch.from_numpy(D).reshape(-1).cuda())
        torch.cuda.synchronize()
        l, r = imageL.copy(), imageR.copy()
        assert F.RunFastACVNet_plus_RectifyImage(p, l, r, pointcloud, disparity) == 0
        assert seen[-1] != 0, "the engine callback must be given the handle's stream, not NULL"
        assert np.array_equal(l, rl) and np.array_equal(r, rr)
        assert np.array_equal(disparity.array, D)
        want, _ = O.reproject_dense(O.FASTACVNET_PLUS, D, rl, c["Q"])
        _check_points(pointcloud.reshape(-1, 6), want)
        torch.cuda.synchronize()
        assert np.array_equal(snap[0].cpu().numpy().reshape(3, H, W), tl) and np.array_equal(snap[1].cpu().numpy().reshape(3, H, W), tr)
    # an engine that fails: Run returns -1, the next Run without the engine works again
    assert F.SetEngine(p, lambda *a: 7) == 0
    assert F.RunFastACVNet_plus_RectifyImage(p, imageL.copy(), imageR.copy(), pointcloud, disparity) == -1
    assert "engine" in F.LastError(p)
    assert F.SetEngine(p, None) == 0
    assert F.RunFastACVNet_plus_RectifyImage(p, imageL.copy(), imageR.copy(), pointcloud, disparity) == 0
    assert F.Release(p) == 0


def test_device_resident_step_can_be_captured_in_a_cuda_graph_and_replayed(calibs):
    """Stage A -> (network slot: here the stub) -> stage B -> stage C on device buffers, on a caller's stream, neither allocate nor
    synchronise nor record events (unless SA_FLAG_TIMING asks for them): the whole step can be captured once in a CUDA graph and
    replayed on new frames in the same buffers - the way a serving loop removes the per-launch cost of a batch-1 step.
    (The tensor-copy descriptors of the warp-specialised stage A kernel are kernel parameters: captured by value.)"""
    import torch
    from oracle import oracle as O
    from oracle import oracle_c as OC
    from stereoalgorithms_b200.geometry import DENSE, StereoGeometry
    from test_gpu_parity import assert_close_points
    W, H, n = 640, 480, 3
    c = calibs["calib_a"]
    rng = np.random.default_rng(3)
    frames = [rng.integers(0, 256, (2, n, H, W, 3), dtype=np.uint8) for _ in range(2)]
    dL = torch.empty((n, H, W, 3), dtype=torch.uint8, device="cuda")
    dR = torch.empty_like(dL)
    pts = torch.zeros((n, W * H, 6), dtype=torch.float32, device="cuda")
    heat = torch.zeros((n, H, W, 3), dtype=torch.uint8, device="cuda")
    (l1, l2), (r1, r2) = O.rectify_maps(c, (W, H))
    with StereoGeometry(W, H, model="FastACVNet_plus", max_batch=n) as g:
        g.set_calibration(c)
        s = torch.cuda.Stream()
        sp = s.cuda_stream

        def step():
            g.rectify_preprocess(dL, dR, n=n, stream=sp)
            g.stub_disparity(n, seed=11, stream=sp)
            g.reproject(n, DENSE, out=pts, to_host=False, stream=sp)
            g.heatmap(n, out=heat, to_host=False, stream=sp)

        dL.copy_(torch.from_numpy(frames[0][0])); dR.copy_(torch.from_numpy(frames[0][1]))
        torch.cuda.synchronize()
        with torch.cuda.stream(s):
            step()                                             # warm-up outside the capture (lazy one-time set-up)
        s.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=s):
            step()
        for k in (1, 0):
            dL.copy_(torch.from_numpy(frames[k][0])); dR.copy_(torch.from_numpy(frames[k][1]))
            pts.zero_(); heat.zero_()
            torch.cuda.synchronize()
            graph.replay()
            torch.cuda.synchronize()
            disp = g.disparity(n)
            got, hm = pts.cpu().numpy(), heat.cpu().numpy()
            for i in range(n):
                wl = OC.remap(frames[k][0][i], l1, l2)
                assert np.array_equal(g.rectified(n, 0)[i], wl) and np.array_equal(g.rectified(n, 1)[i], OC.remap(frames[k][1][i], r1, r2)), (k, i)
                want, _ = O.reproject_dense(O.FASTACVNET_PLUS, disp[i], wl, c["Q"])
                assert_close_points(got[i], want, f"graph replay {k}, frame {i}")
                assert np.array_equal(hm[i], O.heatmap(disp[i])), (k, i)
