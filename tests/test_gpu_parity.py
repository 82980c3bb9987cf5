"""Parity of the CUDA path (through the C ABI) with the oracle and the golden vectors.  Needs a B200: -m gpu.

Gates: maps, rectified u8, resized u8, heatmap, f16 tensors, point count and order: EXACT.
f32 tensors and XYZ: |a-b| <= 1e-5 * max(|a|,|b|), inf/NaN equal in kind and position.
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN, md5, scaled_case
from oracle import oracle as O
from oracle import oracle_c as OC

pytestmark = pytest.mark.gpu

RTOL = 1e-5  # north_star: "normalised tensors and XYZ must fall within 1e-5 relative"


def geom(*a, **k):
    from stereoalgorithms_b200.geometry import StereoGeometry
    return StereoGeometry(*a, **k)


def assert_close_points(a, b, tag=""):
    """1e-5 relative; inf/NaN must match in kind and position."""
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape, (tag, a.shape, b.shape)
    assert np.array_equal(np.isnan(a), np.isnan(b)), tag + ": NaN positions differ"
    assert np.array_equal(np.isposinf(a), np.isposinf(b)) and np.array_equal(np.isneginf(a), np.isneginf(b)), tag + ": inf positions differ"
    fin = np.isfinite(b)
    err = np.abs(a[fin] - b[fin])
    tol = RTOL * np.maximum(np.abs(a[fin]), np.abs(b[fin]))
    assert np.all(err <= tol), f"{tag}: max rel err {np.max(err / np.maximum(tol / RTOL, 1e-30))}"


EYE_KEYS = (("intrinsic_left", "distCoeffs_left", "R_L", "P1"), ("intrinsic_right", "distCoeffs_right", "R_R", "P2"))


# ---- a2: maps -----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["calib_a", "calib_b", "calib_c_rational"])
def test_maps_demo_bit_exact(golden, calibs, name):
    g_ = golden["cases"]["demo_" + name]
    with geom(640, 480) as g:
        g.load_calibration_yaml(os.path.join(GOLDEN, name + ".yml"))
        for eye, tag in ((0, "left"), (1, "right")):
            m1, m2 = g.get_maps(eye)
            assert md5(m1) == g_["map1_" + tag] and md5(m2) == g_["map2_" + tag]
            k = EYE_KEYS[eye]
            o1, o2 = O.init_undistort_rectify_map(*(calibs[name][x] for x in k), (640, 480))
            assert np.array_equal(m1, o1) and np.array_equal(m2, o2)


@pytest.mark.parametrize("name", ["kitti_1242x375", "hd_1920x1080", "hd_c_1920x1080", "uhd_3840x2160"])
def test_maps_scaled_bit_exact(golden, calibs, name):
    g_ = golden["cases"]["scaled_" + name]
    c, (W, H) = scaled_case(calibs, name)
    with geom(W, H) as g:
        g.set_calibration(c)
        for eye, tag in ((0, "left"), (1, "right")):
            m1, m2 = g.get_maps(eye)
            assert md5(m1) == g_["map1_" + tag] and md5(m2) == g_["map2_" + tag], (name, tag)


@pytest.mark.parametrize("tag", ["absurd_fscale_0.02", "absurd_fscale_1e-7", "absurd_k1_+50"])
def test_maps_absurd_calibrations(golden, calibs, tag):
    g_ = golden["cases"][tag]
    c = {k: np.array(v, copy=True) for k, v in calibs["calib_a"].items()}
    c["P1"][0, 0] *= g_["fscale"]
    c["P1"][1, 1] *= g_["fscale"]
    c["distCoeffs_left"][0, 0] += g_["dk1"]
    img = O.synthetic_image(640, 480, g_["image_seed"])
    with geom(640, 480) as g:
        g.set_calibration(c)
        m1, m2 = g.get_maps(0)
        assert md5(m1) == g_["map1"] and md5(m2) == g_["map2"]
        # and the remap of a frame whose taps are (almost) all outside the image
        g.rectify_preprocess(img[None], img[None])
        assert md5(g.rectified(1, 0)[0]) == g_["rect"]


# ---- a3: rectified images -------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["calib_a", "calib_b", "calib_c_rational"])
@pytest.mark.parametrize("variant", ["default", "tile_fastest", "force_direct", "force_generic", "persistent", "persistent_tile_fastest", "per_tile_cta", "no_bulk", "force_unaligned",
                                     "no_l2_prefetch", "no_fixed_size", "no_align16", "no_align16_no_fixed_size"])
def test_rectified_demo_bit_exact(golden, demo_pair, name, variant):
    g_ = golden["cases"]["demo_" + name]
    l, r = demo_pair
    with geom(640, 480, model="RAFTStereo") as g:
        if variant == "force_direct":
            g.set_option("force_direct", 1)
        if variant.endswith("tile_fastest"):
            g.set_option("frame_fastest", 0)
        if variant.startswith("persistent"):
            g.set_option("batch_loop", 0)
            g.set_option("persistent", 4)
        if variant in ("per_tile_cta", "tile_fastest", "force_generic"):
            g.set_option("batch_loop", 0)
        if variant == "no_bulk":
            g.set_option("bulk", 0)
        if variant == "force_unaligned":
            g.set_option("force_unaligned", 1)       # the W % 4 != 0 code path (funnel-shifted loads, run-wise word stores) on aligned frames
        if variant == "no_l2_prefetch":
            g.set_option("l2_prefetch", 0)
        if variant.startswith("no_align16"):
            g.set_option("no_align16", 1)             # the 4-byte-aligned frame loop where the 16-byte one would apply
        if variant.endswith("no_fixed_size"):
            g.set_option("no_fixed_size", 1)          # the run-time-size instantiation of the batch kernel at 640x480
        if variant == "force_generic":
            g.set_option("force_generic", 1)
        g.load_calibration_yaml(os.path.join(GOLDEN, name + ".yml"))
        g.rectify_preprocess(l[None], r[None])
        assert md5(g.rectified(1, 0)[0]) == g_["rect_left"]
        assert md5(g.rectified(1, 1)[0]) == g_["rect_right"]


@pytest.mark.parametrize("name", ["kitti_1242x375", "hd_1920x1080", "hd_c_1920x1080", "uhd_3840x2160"])
def test_rectified_scaled_bit_exact(golden, calibs, name):
    g_ = golden["cases"]["scaled_" + name]
    c, (W, H) = scaled_case(calibs, name)
    li, ri = O.synthetic_image(W, H, g_["image_seeds"][0]), O.synthetic_image(W, H, g_["image_seeds"][1])
    for batch_loop, persistent, bulk, no16 in ((1, 0, 1, 0), (1, 0, 0, 0), (1, 0, 0, 1), (0, 4, 0, 0), (0, 0, 0, 0)):
        with geom(W, H, model="CREStereo") as g:
            g.set_option("no_align16", no16)
            g.set_option("batch_loop", batch_loop)
            g.set_option("persistent", persistent)
            g.set_option("bulk", bulk)
            g.set_calibration(c)
            g.rectify_preprocess(li[None], ri[None])
            assert md5(g.rectified(1, 0)[0]) == g_["rect_left"], name
            assert md5(g.rectified(1, 1)[0]) == g_["rect_right"], name


@pytest.mark.parametrize("model,dtype", [("FastACVNet_plus", "f32"), ("RAFTStereo", "f16"), ("HitNet", "f16"), ("CREStereo", "f32")])
@pytest.mark.parametrize("size", ["hd_1920x1080", "sd_1280x720"])
def test_batches_on_the_warp_specialised_kernel(calibs, model, dtype, size):
    """rectify_ws_kernel (tensor copies, producer / consumer warps, two- and three-stage rings) on sizes with partial bottom
    tiles: 1920x1080 (its own compile-time instantiation; 1080 = 67.5 tiles) and 1280x720 (run-time size), frame counts that
    wrap the rings unevenly, all four normalisations: rectified images and tensors exact against the oracle."""
    from stereoalgorithms_b200.geometry import F16, F32, MODELS
    if size == "hd_1920x1080":
        c, (W, H) = scaled_case(calibs, size)
    else:
        W, H = 1280, 720
        c = O.scale_calibration(calibs["calib_b"], 2.0, crop_y=(480 * 2.0 - H) / 2)
    n = 5
    L = np.stack([O.synthetic_image(W, H, 400 + i) for i in range(n)])
    R = np.stack([O.synthetic_image(W, H, 500 + i) for i in range(n)])
    (l1, l2), (r1, r2) = O.rectify_maps(c, (W, H))
    npdt = np.float16 if dtype == "f16" else np.float32
    with geom(W, H, model=model, dtype=F16 if dtype == "f16" else F32, max_batch=n) as g:
        g.set_calibration(c)
        g.set_option("batch_ctas", 1)                        # the whole batch in one CTA per tile: 5 frames = 2 unrolled pairs + 1
        g.rectify_preprocess(L, R)
        gl, gr = g.rectified(n, 0), g.rectified(n, 1)
        t = g.tensor(n) if model == "HitNet" else (g.tensor(n, 0), g.tensor(n, 1))
        for i in range(n):
            wl, wr = OC.remap(L[i], l1, l2), OC.remap(R[i], r1, r2)
            assert np.array_equal(gl[i], wl) and np.array_equal(gr[i], wr), i
            want = O.preprocess(MODELS[model], wl, wr, npdt)
            if model == "HitNet":
                assert np.array_equal(t[i], want), i
            else:
                assert np.array_equal(t[0][i], want[0]) and np.array_equal(t[1][i], want[1]), i
        g.set_option("batch_ctas", 1000)                     # one frame per CTA
        g.rectify_preprocess(L[:2], R[:2])
        assert np.array_equal(g.rectified(2, 1), gr[:2])


def test_rectified_batch_device_input_and_pitch(calibs, demo_pair):
    """Batch of different frames, device-resident input (torch), then host input with a row pitch."""
    import torch
    c = calibs["calib_b"]
    n = 5
    rng = np.random.default_rng(3)
    L = np.stack([np.clip(demo_pair[0].astype(np.int16) + rng.integers(-8, 9, demo_pair[0].shape), 0, 255).astype(np.uint8) for _ in range(n)])
    R = np.stack([np.clip(demo_pair[1].astype(np.int16) + rng.integers(-8, 9, demo_pair[1].shape), 0, 255).astype(np.uint8) for _ in range(n)])
    (l1, l2), (r1, r2) = O.rectify_maps(c, (640, 480))
    want_l = np.stack([OC.remap(x, l1, l2) for x in L])
    want_r = np.stack([OC.remap(x, r1, r2) for x in R])
    with geom(640, 480, model="HitNet", max_batch=8) as g:
        g.set_calibration(c)
        g.rectify_preprocess(torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda())
        assert np.array_equal(g.rectified(n, 0), want_l) and np.array_equal(g.rectified(n, 1), want_r)
        # pitched host rows (a cv::Mat ROI): 1920 -> 2000 bytes per row
        Lp = np.zeros((n, 480, 2000), np.uint8)
        Rp = np.zeros((n, 480, 2000), np.uint8)
        Lp[:, :, :1920] = L.reshape(n, 480, 1920)
        Rp[:, :, :1920] = R.reshape(n, 480, 1920)
        g.rectify_preprocess(Lp, Rp, n=n, pitch=2000)
        assert np.array_equal(g.rectified(n, 0), want_l) and np.array_equal(g.rectified(n, 1), want_r)


def test_the_benchmarked_configuration_itself(calibs, demo_pair):
    """BASELINE configs[1] exactly as bench.py runs it: 64 pairs of 640x480, RAFTStereo, f16 tensors, every option at its
    default, device-resident input, stage A then stage B back to back on a caller-owned stream (the multi-chunk, tapered
    schedule a 64-pair batch gets; the <F16, IDENT, 640, 480> instantiation).  A sample of the frames - the first and last
    frame of every chunk region among them - is checked against the oracle: rectified images and f16 tensors exactly, the
    flipped disparity exactly, XYZ within the gate."""
    import torch
    from stereoalgorithms_b200.geometry import DENSE, F16
    c = calibs["calib_a"]
    n, W, H = 64, 640, 480
    L = np.empty((n, H, W, 3), np.uint8)
    R = np.empty((n, H, W, 3), np.uint8)
    for i in range(n):                                        # bench.py:make_inputs - the demo pair plus +-8 noise, seed = frame
        rng = np.random.default_rng(i)
        L[i] = np.clip(demo_pair[0].astype(np.int16) + rng.integers(-8, 9, demo_pair[0].shape), 0, 255)
        R[i] = np.clip(demo_pair[1].astype(np.int16) + rng.integers(-8, 9, demo_pair[1].shape), 0, 255)
    D = np.stack([-O.synthetic_disparity(W, H, 900 + i) for i in range(n)])    # the RAFT network emits the negated disparity
    stream = torch.cuda.Stream()
    with geom(W, H, model="RAFTStereo", dtype=F16, max_batch=n) as g:
        g.set_calibration(c)
        with torch.cuda.stream(stream):
            dL, dR = torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()
            g.set_disparity(torch.from_numpy(D[:, None]).cuda(), n=n, stream=stream.cuda_stream)
            for _ in range(2):                                # twice: the sign flip is written back, so flip it back in between
                g.rectify_preprocess(dL, dR, n=n, stream=stream.cuda_stream)
                g.reproject(n, DENSE, to_host=False, stream=stream.cuda_stream)
            g.reproject(n, DENSE, to_host=False, stream=stream.cuda_stream)   # odd number of flips in total: disparity is +d now
        stream.synchronize()
        rl, rr = g.rectified(n, 0), g.rectified(n, 1)
        tl, tr = g.tensor(n, 0), g.tensor(n, 1)
        disp = g.disparity(n)
        pts = np.empty((n, W * H, 6), np.float32)
        g.download(n, pointcloud=pts)
    assert np.array_equal(disp, -D)
    (l1, l2), (r1, r2) = O.rectify_maps(c, (W, H))
    for i in (0, 1, 11, 12, 13, 26, 27, 31, 32, 40, 52, 62, 63):
        wl, wr = OC.remap(L[i], l1, l2), OC.remap(R[i], r1, r2)
        assert np.array_equal(rl[i], wl) and np.array_equal(rr[i], wr), i
        a, b = O.preprocess(O.RAFTSTEREO, wl, wr, np.float16)
        assert np.array_equal(tl[i], a) and np.array_equal(tr[i], b), i
        want, _ = O.reproject_dense(O.RAFTSTEREO, D[i], wl, c["Q"])
        assert_close_points(pts[i], want, f"bench configuration, frame {i}")


# ---- a4 / a10: tensors -----------------------------------------------------------------------------------------
@pytest.mark.parametrize("model", ["RAFTStereo", "HitNet", "CREStereo", "FastACVNet_plus"])
@pytest.mark.parametrize("dtype", ["f32", "f16"])
@pytest.mark.parametrize("path", ["rectify", "preprocess4", "preprocess1"])
def test_tensors(calibs, demo_pair, model, dtype, path):
    """path: fused rectify + preprocess; rectify = 0 with the four-pixel-per-thread kernel; rectify = 0 with the one-pixel one."""
    from stereoalgorithms_b200.geometry import F16, F32, MODELS
    c = calibs["calib_a"]
    l, r = demo_pair
    rectify = path == "rectify"
    npdt = np.float16 if dtype == "f16" else np.float32
    rl, rr = O.rectify_pair(c, l, r) if rectify else (l, r)
    want = O.preprocess(MODELS[model], rl, rr, npdt)
    with geom(640, 480, model=model, dtype=F16 if dtype == "f16" else F32, max_batch=2) as g:
        g.set_option("scalar_preprocess", 1 if path == "preprocess1" else 0)
        g.set_calibration(c)
        g.rectify_preprocess(np.stack([l, l]), np.stack([r, r]), rectify=rectify)
        if model == "HitNet":
            t = g.tensor(2)
            assert t.shape == (2, 6, 480, 640) and t.dtype == npdt
            got = [t[0], t[1]]
            want = [want, want]
        else:
            tl, tr = g.tensor(2, 0), g.tensor(2, 1)
            got = [tl[0], tr[0], tl[1], tr[1]]
            want = [want[0], want[1], want[0], want[1]]
        for a, b in zip(got, want):
            if dtype == "f16":
                assert np.array_equal(a, b)                       # gate: exact
            else:
                assert np.all(np.abs(a - b) <= RTOL * np.maximum(np.abs(a), np.abs(b)))   # gate: 1e-5 relative
                assert np.array_equal(a, b)                       # and in fact identical (LUT built the same way)


@pytest.mark.parametrize("size,pitch", [((644, 37), 0), ((640, 33), 1936), ((322, 41), 0), ((64, 3), 0), ((333, 77), 0), ((1242, 375), 0),
                                        ((321, 5), 970), ((3, 1), 0)])
@pytest.mark.parametrize("model,dtype", [("FastACVNet_plus", "f32"), ("HitNet", "f16"), ("RAFTStereo", "f16"), ("CREStereo", "f32")])
def test_tensors_without_rectification_odd_shapes(size, pitch, model, dtype):
    """rectify = 0 on device input.  Three kernels share this path: four pixels per thread for word-aligned rows of W % 4 = 0
    (644 = 161 groups per row: partial last CTA; pitched rows), the linear four-pixel kernel for dense frames of any width
    (planes aligned to 8, 4 or 2 bytes only; npix % 4 tail pixels), one pixel per thread for the rest (odd width AND pitch)."""
    import torch
    from stereoalgorithms_b200.geometry import F16, F32, MODELS
    W, H = size
    n = 3
    rng = np.random.default_rng(W * H)
    L = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    R = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    if pitch:
        Lp, Rp = np.zeros((n, H, pitch), np.uint8), np.zeros((n, H, pitch), np.uint8)
        Lp[:, :, :3 * W], Rp[:, :, :3 * W] = L.reshape(n, H, -1), R.reshape(n, H, -1)
        dl, dr = torch.from_numpy(Lp).cuda(), torch.from_numpy(Rp).cuda()
    else:
        dl, dr = torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()
    npdt = np.float16 if dtype == "f16" else np.float32
    with geom(W, H, model=model, dtype=F16 if dtype == "f16" else F32, max_batch=n) as g:
        g.rectify_preprocess(dl, dr, n=n, pitch=pitch, rectify=False)
        tl = g.tensor(n, 0)
        tr = None if model == "HitNet" else g.tensor(n, 1)
        for i in range(n):
            want = O.preprocess(MODELS[model], L[i], R[i], npdt)
            if model == "HitNet":
                assert np.array_equal(tl[i], want), (size, i)
            else:
                assert np.array_equal(tl[i], want[0]) and np.array_equal(tr[i], want[1]), (size, i)


# ---- a5 / a9: reprojection ---------------------------------------------------------------------------------------
def _disp_input(model_id, W, H, seed):
    d = O.synthetic_disparity(W, H, seed)
    if model_id == O.RAFTSTEREO:
        d = -d                                           # the network emits negative flow; the kernel flips it
    if model_id == O.CRESTEREO:
        return np.stack([d, np.full_like(d, 7.0)])       # 2 planes, plane 0 is the disparity
    return d[None]


@pytest.mark.parametrize("model", ["RAFTStereo", "HitNet", "CREStereo", "FastACVNet_plus"])
@pytest.mark.parametrize("calib", ["calib_a", "calib_c_rational"])
@pytest.mark.parametrize("exact_div", [0, 1])
def test_reproject_dense_and_compact(calibs, demo_pair, model, calib, exact_div):
    """exact_div = 1: IEEE divisions, results bit-identical to the oracle.  exact_div = 0 (default): the three quotients
    of a pixel share one refined reciprocal - within 1 ulp, far inside the 1e-5 gate; inf/NaN identical in kind/position."""
    from stereoalgorithms_b200.geometry import COMPACT, DENSE, MODELS
    mid = MODELS[model]
    c = calibs[calib]
    l, r = demo_pair
    n = 3
    L, R = np.stack([l] * n), np.stack([r] * n)
    D = np.stack([_disp_input(mid, 640, 480, 100 + i) for i in range(n)])
    rl, _ = O.rectify_pair(c, l, r)
    with geom(640, 480, model=model, max_batch=4) as g:
        g.set_calibration(c)
        g.set_option("exact_div", exact_div)
        g.rectify_preprocess(L, R)
        g.set_disparity(D)
        pts, cnt = g.reproject(n, DENSE)
        dout = g.disparity(n)
        assert list(cnt) == [640 * 480] * n
        for i in range(n):
            want, dw = O.reproject_dense(mid, D[i, 0], rl, c["Q"])
            assert_close_points(pts[i], want, f"{model} dense frame {i}")
            assert np.array_equal(pts[i][:, 3:], want[:, 3:])                 # colours exact
            if exact_div:
                assert np.array_equal(pts[i].view(np.uint32), want.view(np.uint32))
            else:                                                             # <= 1 ulp in practice
                fin = np.isfinite(want[:, :3]) & (want[:, :3] != 0)
                ulp = np.abs(pts[i][:, :3].view(np.int32).astype(np.int64) - want[:, :3].view(np.int32).astype(np.int64))
                assert ulp[fin].max() <= 1
            assert np.array_equal(dout[i], dw)                                # RAFT: flipped and written back
        # compact (extension): count and order exact
        g.set_disparity(D)
        cp, cc = g.reproject(n, COMPACT)
        for i in range(n):
            want, k, _ = O.reproject_compact(mid, D[i, 0], rl, c["Q"])
            assert cc[i] == k and 0 < k < 640 * 480
            assert_close_points(cp[i, :k], want, f"{model} compact frame {i}")
            assert np.array_equal(cp[i, :k, 3:], want[:, 3:])                   # same pixels in the same order (colours exact)
            if exact_div:
                assert np.array_equal(cp[i, :k].view(np.uint32), want.view(np.uint32))
            assert not cp[i, k:].any()                                          # nothing written past the count


@pytest.mark.parametrize("size", [(1242, 375), (322, 241), (64, 16), (5, 3)])
def test_reproject_odd_sizes(calibs, size):
    """Pixel counts that are not multiples of 4 / 1024: scalar paths, partial warps, rows that straddle a thread."""
    from stereoalgorithms_b200.geometry import COMPACT, DENSE
    W, H = size
    c = O.scale_calibration(calibs["calib_a"], W / 640)
    rng = np.random.default_rng(W * H)
    img = rng.integers(0, 256, (2, H, W, 3), dtype=np.uint8)
    d = rng.normal(20, 15, (2, 1, H, W)).astype(np.float32)
    d[rng.random(d.shape) < 0.1] = 0
    with geom(W, H, model="FastACVNet_plus", max_batch=2) as g:
        g.set_calibration(c)
        g.rectify_preprocess(img, img, rectify=False)
        g.set_disparity(d)
        pts, _ = g.reproject(2, DENSE)
        cp, cc = g.reproject(2, COMPACT)                   # count -> scan -> compact, three launches (default)
        g.set_option("single_pass_compact", 1)
        cp2, cc2 = g.reproject(2, COMPACT)                 # single-pass (decoupled look-back) kernel
        assert np.array_equal(cc, cc2) and np.array_equal(cp, cp2, equal_nan=True)
        heat = g.heatmap(2)
        for i in range(2):
            want, _ = O.reproject_dense(O.FASTACVNET_PLUS, d[i, 0], img[i], c["Q"])
            assert_close_points(pts[i], want, f"{size} dense")
            wc, k, _ = O.reproject_compact(O.FASTACVNET_PLUS, d[i, 0], img[i], c["Q"])
            assert cc[i] == k
            assert_close_points(cp[i, :k], wc, f"{size} compact")
            assert np.array_equal(cp[i, :k, 3:], wc[:, 3:])
            assert np.array_equal(heat[i], O.heatmap(d[i, 0]))


def test_compaction_single_pass_long_frames_many_launches(calibs):
    """The single-pass compaction (option single_pass_compact) on frames of 2025 tiles (1920x1080: look-back windows of 32 tiles, many tiles in flight), a batch
    that is launched several times in a row (the look-back words are never cleared: only the launch's epoch makes them valid),
    with sparse and dense validity: counts and order exact against the oracle every time."""
    from stereoalgorithms_b200.geometry import COMPACT
    W, H, n = 1920, 1080, 3
    c = O.scale_calibration(calibs["calib_a"], 3.0)
    rng = np.random.default_rng(11)
    img = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    with geom(W, H, model="FastACVNet_plus", max_batch=n) as g:
        g.set_calibration(c)
        g.set_option("single_pass_compact", 1)
        g.rectify_preprocess(img, img, rectify=False)
        for density in (0.5, 0.02, 1.0, 0.0):
            d = rng.uniform(1, 60, (n, 1, H, W)).astype(np.float32)
            d[rng.random(d.shape) >= density] = -1.0
            g.set_disparity(d)
            cp, cc = g.reproject(n, COMPACT)
            for i in range(n):
                wc, k, _ = O.reproject_compact(O.FASTACVNET_PLUS, d[i, 0], img[i], c["Q"])
                assert cc[i] == k, (density, i)
                if k:
                    assert np.array_equal(cp[i, :k, 3:], wc[:, 3:]), (density, i)        # same pixels in the same order
                    assert_close_points(cp[i, :k:97], wc[::97], f"single-pass compact, density {density}")


@pytest.mark.parametrize("size", [(1242, 375), (638, 50), (333, 77), (67, 19), (9, 7)])
@pytest.mark.parametrize("source", ["rectified", "pitched_input"])
def test_reproject_pitched_colour_rows(calibs, size, source):
    """Stage B reading colours from rows that are NOT contiguous: the internal rectified buffer of widths that are not a
    multiple of 4 (rows padded to whole 4-pixel groups), and a caller's pitched image (rectify = 0).  Pixel groups that run
    over a row end, the last groups of a frame (no slack after the last row) and tiny images take different code paths."""
    from stereoalgorithms_b200.geometry import COMPACT, DENSE
    W, H = size
    s = W / 640
    c = O.scale_calibration(calibs["calib_a"], s, crop_y=(480 * s - H) / 2)
    n = 3
    rng = np.random.default_rng(W + 7 * H)
    L = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    d = rng.normal(20, 15, (n, 1, H, W)).astype(np.float32)
    d[rng.random(d.shape) < 0.1] = 0
    with geom(W, H, model="FastACVNet_plus", max_batch=n) as g:
        g.set_calibration(c)
        if source == "rectified":
            g.rectify_preprocess(L, L[::-1].copy())
            (l1, l2), _ = O.rectify_maps(c, (W, H))
            colour = np.stack([OC.remap(x, l1, l2) for x in L])
        else:
            pitch = 3 * W + 5                                           # odd pitch: rows at every byte alignment
            Lp = np.zeros((n, H, pitch), np.uint8)
            Lp[:, :, :3 * W] = L.reshape(n, H, 3 * W)
            Lp = Lp.reshape(-1)[: (n - 1) * H * pitch + (H - 1) * pitch + 3 * W].copy()   # no slack after the last row
            import torch
            Ld = torch.from_numpy(Lp).cuda()                            # device input is read in place, with its pitch
            g.rectify_preprocess(Ld, Ld, n=n, pitch=pitch, rectify=False)
            colour = L
        g.set_disparity(d)
        pts, _ = g.reproject(n, DENSE)
        cp, cc = g.reproject(n, COMPACT)
        for i in range(n):
            want, _ = O.reproject_dense(O.FASTACVNET_PLUS, d[i, 0], colour[i], c["Q"])
            assert_close_points(pts[i], want, f"{size} {source} dense")
            assert np.array_equal(pts[i][:, 3:], want[:, 3:])
            wc, k, _ = O.reproject_compact(O.FASTACVNET_PLUS, d[i, 0], colour[i], c["Q"])
            assert cc[i] == k
            assert np.array_equal(cp[i, :k, 3:], wc[:, 3:])


def test_reproject_general_Q(calibs, demo_pair):
    """A Q with every used entry non-trivial: products are no longer exact, hence the 1e-5 gate."""
    from stereoalgorithms_b200.geometry import DENSE
    c = calibs["calib_a"]
    Q = np.array([[1.01, 0.02, 0, -295.1], [-0.03, 0.99, 0, -274.7], [0.001, -0.002, 0, 541.6], [0, 0, 0.01663, 0.1]])
    d = O.synthetic_disparity(640, 480, 5)
    with geom(640, 480) as g:
        g.set_calibration(c)
        g.set_Q(Q)
        g.rectify_preprocess(demo_pair[0][None], demo_pair[1][None], rectify=False)
        g.set_disparity(d[None, None])
        pts, _ = g.reproject(1, DENSE)
        want, _ = O.reproject_dense(O.FASTACVNET_PLUS, d, demo_pair[0], Q)
        assert_close_points(pts[0], want, "general Q")
        assert np.array_equal(g.get_Q(), Q)


# ---- a6: heatmap ---------------------------------------------------------------------------------------------------
def test_heatmap_golden(golden):
    cases = {"synthetic_640x480": O.synthetic_disparity(640, 480, 1234),
             "synthetic_noinvalid": O.synthetic_disparity(640, 480, 1, invalid=False),
             "negated": -O.synthetic_disparity(320, 240, 5),
             "constant": np.full((48, 64), 3.25, np.float32),
             "random": np.random.default_rng(9).normal(30, 20, (375, 1242)).astype(np.float32)}
    for tag, d in cases.items():
        H, W = d.shape
        with geom(W, H, model="HitNet") as g:
            g.set_disparity(d[None, None])
            hm = g.heatmap(1)                      # standalone min/max pass
            assert md5(hm[0]) == golden["heatmap"][tag]["md5"], tag


@pytest.mark.parametrize("fold", [False, True])
def test_heatmap_ties_and_near_ties(calibs, fold):
    """The heat map index is round_half_even(|fl(fl(d - min) / s)|).  The kernel multiplies by 1/s and takes the integer from
    the mantissa of q + 512.5..., falling back to the IEEE division only within 2/16384 of a half-integer: disparities that
    land exactly on half-integers (s = 1, 1/2, 2: ties, half of them round DOWN to even), one and a few ulps either side of them,
    and on arbitrary scales must come out as the reference's arithmetic gives them.  fold: through stage B's min/max partials."""
    from stereoalgorithms_b200.geometry import DENSE
    W, H = 640, 480
    rng = np.random.default_rng(77)
    frames = []
    for mn, mx in ((0.0, 255.0), (10.0, 137.5), (-3.0, 507.0), (1.25, 77.0), (-40.0, 3.1), (100.0, 100.0 + 255 * 0.3)):
        mn32, mx32 = np.float32(mn), np.float32(mx)
        s = np.float32((float(mx32) - float(mn32)) / 255)
        k = rng.integers(0, 255, W * H).astype(np.float64) + 0.5
        d = (np.float64(mn32) + k * np.float64(s)).astype(np.float32)
        ulps = rng.integers(-3, 4, W * H)
        for u in range(-3, 4):
            m = ulps == u
            t = d[m]
            for _ in range(abs(u)):
                t = np.nextafter(t, np.float32(np.inf if u > 0 else -np.inf))
            d[m] = t
        d[:1000] = rng.uniform(mn, mx, 1000).astype(np.float32)
        d = np.clip(d, mn32, mx32)
        d[0], d[1] = mn32, mx32
        frames.append(d.reshape(H, W))
    d = np.stack(frames)
    n = len(frames)
    with geom(W, H, model="FastACVNet_plus", max_batch=n) as g:
        g.set_calibration(calibs["calib_a"])
        g.set_disparity(d[:, None])
        if fold:
            img = np.zeros((n, H, W, 3), np.uint8)
            g.rectify_preprocess(img, img)
            g.reproject(n, DENSE, to_host=False)
        hm = g.heatmap(n)
    for i in range(n):
        want = O.heatmap(d[i])
        assert np.array_equal(hm[i], want), (i, int((hm[i] != want).any(axis=2).sum()))


def test_disparity_u8_golden(golden):
    """sa_disparity_u8 = the convertTo(CV_8U) inside cv::imwrite("disparity.jpg", disparity): vectors made by cv2."""
    import torch
    cases = {"special": np.load(os.path.join(GOLDEN, "disparity_u8_special.npy")),
             "synthetic_640x480": O.synthetic_disparity(640, 480, 1234),
             "negated": -O.synthetic_disparity(320, 240, 5),
             "random": np.random.default_rng(9).normal(30, 80, (375, 1242)).astype(np.float32)}
    for tag, d in cases.items():
        H, W = d.shape
        with geom(W, H, model="CREStereo", max_batch=2) as g:          # two planes per frame: plane 0 is the disparity
            x = np.full((2, 2, H, W), 777.0, np.float32)                # plane 1 must be ignored
            x[0, 0], x[1, 0] = d, d[::-1]
            g.set_disparity(x)
            img = g.disparity_u8(2)
            assert md5(img[0]) == golden["disparity_u8"][tag]["md5"], tag
            assert np.array_equal(img[1], O.disparity_u8(d[::-1])), tag
            dev = torch.empty((2, H, W), dtype=torch.uint8, device="cuda")
            g.disparity_u8(2, out=dev, to_host=False)
            g.synchronize()
            assert np.array_equal(dev.cpu().numpy(), img), tag


@pytest.mark.parametrize("W,H,launches", [(640, 480, 2), (1920, 1080, 3)])
def test_heatmap_after_reproject_uses_fused_minmax(calibs, W, H, launches):
    """Stage B leaves per-CTA min/max partials; the heat map kernel folds them itself (up to 1024 per frame: one launch), larger
    frames get the separate fold launch first.  Either way no second pass over the disparity, and a repeated call reuses the pair."""
    from stereoalgorithms_b200.geometry import DENSE
    d = np.stack([O.synthetic_disparity(W, H, 40 + i) * (1 + i) for i in range(3)])
    img = np.stack([O.synthetic_image(W, H, 5 + i) for i in range(3)])
    calib = calibs["calib_a"] if W == 640 else O.scale_calibration(calibs["calib_a"], W / 640, crop_y=(480 * W / 640 - H) / 2)
    with geom(W, H, model="RAFTStereo", max_batch=3) as g:
        g.set_calibration(calib)
        g.rectify_preprocess(img, img)
        g.set_disparity(-d[:, None])
        before = g.launch_count
        g.reproject(3, DENSE, to_host=False)
        hm = g.heatmap(3)
        assert g.launch_count - before == launches
        for i in range(3):
            assert np.array_equal(hm[i], O.heatmap(d[i])), i   # on the flipped (positive) disparity
        before = g.launch_count
        hm2 = g.heatmap(3)                                     # the frames' final min/max pairs are current now
        assert g.launch_count - before == 1 and np.array_equal(hm2, hm)


# ---- a8: resize path -------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("src,dst", [((1920, 1080), (640, 480)), ((1280, 960), (640, 480)), ((1242, 375), (640, 192)), ((1242, 375), (318, 96)),
                                     ((1918, 1078), (959, 539)), ((4000, 600), (402, 60))])
@pytest.mark.parametrize("model,dtype", [("FastACVNet_plus", "f32"), ("RAFTStereo", "f16")])
@pytest.mark.parametrize("fused", [1, 0, -1])
def test_resize_path(calibs, src, dst, model, dtype, fused):
    """a8: rectify at full size -> OpenCV's 8-bit INTER_LINEAR downscale (exact 2x: rounded 2x2 mean) -> tensors; output widths
    that are / are not multiples of 4, pitched rectified rows of odd source widths, a ~10x reduction; both tensor types.
    fused = 1 (option fused_resize): the one-launch rectify + resize + normalise kernel where it applies (even widths: host frames
    are uploaded with rows padded to 16 bytes - checked by the launch count); fused = 0, the default: rectify, then resize +
    normalise in one launch (resize_norm_kernel); fused = -1 (option separate_preprocess): rectify, resize, preprocess."""
    from stereoalgorithms_b200.geometry import DENSE, F16, F32, MODELS
    W, H = src
    s = W / 640
    c = O.scale_calibration(calibs["calib_a"], s, crop_y=(480 * s - H) / 2)
    li, ri = O.synthetic_image(W, H, 50), O.synthetic_image(W, H, 51)
    d = O.synthetic_disparity(dst[0], dst[1], 52)
    mid = MODELS[model]
    ref = O.run_pair(mid, c, li, ri, -d if model == "RAFTStereo" else d, net_size=dst, dtype=np.float16 if dtype == "f16" else np.float32)
    with geom(W, H, model=model, net_size=dst, dtype=F16 if dtype == "f16" else F32, max_batch=2) as g:
        g.set_calibration(c)
        g.set_option("fused_resize", 1 if fused == 1 else 0)
        g.set_option("separate_preprocess", 1 if fused == -1 else 0)
        g.set_option("bulk", 2)                               # the tensor-copy kernels for every even width, not only multiples of 16
        k0 = g.launch_count
        g.rectify_preprocess(np.stack([li, ri]), np.stack([ri, li]))
        assert g.launch_count - k0 == (1 if fused == 1 and W % 2 == 0 else (3 if fused == -1 else 2))
        assert np.array_equal(g.rectified(1, 0)[0], ref["rect_left"])
        assert np.array_equal(g.rectified(2, 1)[1], OC.remap(li, *O.rectify_maps(c, (W, H))[1]))
        nl, nr = g.net_image(2, 0), g.net_image(2, 1)
        assert np.array_equal(nl[0], ref["net_left"]) and np.array_equal(nr[0], ref["net_right"])
        tl, tr = g.tensor(2, 0), g.tensor(2, 1)
        assert np.array_equal(tl[0], ref["tensors"][0]) and np.array_equal(tr[0], ref["tensors"][1])
        # second frame of the batch (eyes swapped): resized images / tensors must be those of the swapped rectifications
        (l1, l2), (r1, r2) = O.rectify_maps(c, (W, H))
        assert np.array_equal(nl[1], O.resize_linear_u8(OC.remap(ri, l1, l2), dst)) and np.array_equal(nr[1], O.resize_linear_u8(OC.remap(li, r1, r2), dst))
        g.set_disparity(np.stack([-d if model == "RAFTStereo" else d] * 2)[:, None])
        pts, _ = g.reproject(2, DENSE)
        assert_close_points(pts[0], ref["points"], "resize path")


# ---- a7: the whole Run flow with host buffers ----------------------------------------------------------------------
@pytest.mark.parametrize("model", ["RAFTStereo", "FastACVNet_plus"])
def test_run_batch_host_to_host(calibs, demo_pair, model):
    from stereoalgorithms_b200.geometry import MODELS, PinnedArray
    mid = MODELS[model]
    c = calibs["calib_a"]
    n = 11                                            # not a multiple of the chunk size
    rng = np.random.default_rng(8)
    L = np.stack([np.roll(demo_pair[0], i * 7, axis=1) for i in range(n)])
    R = np.stack([np.roll(demo_pair[1], i * 7, axis=1) for i in range(n)])
    D = np.stack([_disp_input(mid, 640, 480, 200 + i) for i in range(n)])
    pin = [PinnedArray(L.shape, np.uint8), PinnedArray(R.shape, np.uint8), PinnedArray((n, 640 * 480, 6), np.float32),
           PinnedArray((n, 480, 640), np.float32), PinnedArray(L.shape, np.uint8), PinnedArray(R.shape, np.uint8),
           PinnedArray((n, 480, 640, 3), np.uint8)]
    pin[0].array[:] = L
    pin[1].array[:] = R
    with geom(640, 480, model=model, max_batch=16) as g:
        g.set_calibration(c)
        g.set_option("chunk", 4)
        g.set_disparity(D)
        g.run_batch(pin[0].array, pin[1].array, n, rectify=True, stub_seed=0, pointcloud=pin[2].array, disparity=pin[3].array,
                    rect_left=pin[4].array, rect_right=pin[5].array, heatmap=pin[6].array)
        for i in range(n):
            ref = O.run_pair(mid, c, L[i], R[i], D[i, 0])
            assert np.array_equal(pin[4].array[i], ref["rect_left"]) and np.array_equal(pin[5].array[i], ref["rect_right"])
            assert_close_points(pin[2].array[i], ref["points"], f"run_batch {i}")
            assert np.array_equal(pin[3].array[i], ref["disparity"])
            assert np.array_equal(pin[6].array[i], ref["heatmap"])
        # the stub engine produces a usable disparity too (~5 % zeros, ~1 % negatives after the RAFT flip)
        g.run_batch(pin[0].array, pin[1].array, n, rectify=True, stub_seed=7, pointcloud=pin[2].array, disparity=pin[3].array)
        dd = pin[3].array
        assert 0.03 < (dd == 0).mean() < 0.07 and 0.005 < (dd < 0).mean() < 0.015 and dd.max() < 65
        ref_pts, _ = O.reproject_dense(O.FASTACVNET_PLUS, dd[2], O.rectify_pair(c, L[2], R[2])[0], c["Q"])
        assert_close_points(pin[2].array[2], ref_pts, "stub frame 2")


@pytest.mark.parametrize("size", [(1242, 375), (322, 241)])
def test_run_batch_odd_width(calibs, size):
    """The whole host-to-host flow at a width that is not a multiple of 4: the rectified rows are padded on the device, the
    caller still gets dense images back; ordinary (pageable) host buffers."""
    W, H = size
    s = W / 640
    c = O.scale_calibration(calibs["calib_a"], s, crop_y=(480 * s - H) / 2)
    n = 5
    rng = np.random.default_rng(W)
    L = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    R = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    D = np.stack([O.synthetic_disparity(W, H, 300 + i) for i in range(n)])[:, None]
    pts, disp = np.zeros((n, W * H, 6), np.float32), np.zeros((n, H, W), np.float32)
    rl, rr, hm = np.zeros_like(L), np.zeros_like(R), np.zeros((n, H, W, 3), np.uint8)
    with geom(W, H, model="CREStereo", max_batch=8) as g:
        g.set_calibration(c)
        g.set_option("chunk", 2)
        dd = np.zeros((n, 2, H, W), np.float32)
        dd[:, :1] = D
        g.set_disparity(dd)
        g.run_batch(L, R, n, rectify=True, stub_seed=0, pointcloud=pts, disparity=disp, rect_left=rl, rect_right=rr, heatmap=hm)
        for i in range(n):
            ref = O.run_pair(O.CRESTEREO, c, L[i], R[i], D[i, 0])
            assert np.array_equal(rl[i], ref["rect_left"]) and np.array_equal(rr[i], ref["rect_right"]), i
            assert_close_points(pts[i], ref["points"], f"run_batch odd {i}")
            assert np.array_equal(disp[i], ref["disparity"]) and np.array_equal(hm[i], ref["heatmap"])
        assert np.array_equal(g.rectified(n, 1), rr)


# ---- errors ----------------------------------------------------------------------------------------------------------
def test_error_behaviour(demo_pair, tmp_path):
    from stereoalgorithms_b200 import _lib as L
    from stereoalgorithms_b200.geometry import DENSE, StereoGeomError
    l, r = demo_pair
    with geom(640, 480, max_batch=1) as g:
        with pytest.raises(StereoGeomError) as e:
            g.rectify_preprocess(l[None], r[None])                      # no calibration yet
        assert e.value.code == L.SA_ERR_NOT_READY
        with pytest.raises(StereoGeomError) as e:
            g.rectify_preprocess(np.stack([l, l]), np.stack([r, r]))    # n > max_batch
        assert e.value.code == L.SA_ERR_TOO_LARGE
        with pytest.raises(StereoGeomError) as e:
            g.load_calibration_yaml(str(tmp_path / "missing.yml"))
        assert e.value.code == L.SA_ERR_IO
        bad = tmp_path / "bad.yml"
        bad.write_text("%YAML:1.0\nQ: !!opencv-matrix\n   rows: 4\n   cols: 4\n   dt: d\n   data: [ 1., 2. ]\n")
        with pytest.raises(StereoGeomError) as e:
            g.load_calibration_yaml(str(bad))
        assert e.value.code == L.SA_ERR_PARSE
        g.rectify_preprocess(l[None], r[None], rectify=False)
        with pytest.raises(StereoGeomError) as e:
            g.reproject(1, DENSE)                                       # no Q
        assert e.value.code == L.SA_ERR_NOT_READY
        assert g.rectify_preprocess(l[:0][None], r[:0][None], n=0) == 0  # empty batch is a no-op
    with pytest.raises(StereoGeomError):
        geom(640, 480, net_size=(1280, 960))                             # upscaling is not supported
    with pytest.raises(StereoGeomError):
        geom(0, 480)


def test_reproject_special_disparities(calibs):
    """Zeros, negatives, denormals, huge values, inf and NaN in the disparity: kind and position of inf/NaN must match
    the IEEE arithmetic of the reference in both division modes."""
    from stereoalgorithms_b200.geometry import COMPACT, DENSE
    W, H = 64, 16
    c = O.scale_calibration(calibs["calib_a"], 0.1)
    vals = np.array([0.0, -0.0, 1.0, -1.0, 1e-38, -1e-38, 1e-45, 3e38, -3e38, np.inf, -np.inf, np.nan, 1e-30, 1e30, 5.5, 64.0], np.float32)
    d = np.tile(vals, W * H // vals.size).reshape(1, 1, H, W).astype(np.float32)
    img = O.synthetic_image(W, H, 4)
    for exact in (0, 1):
        with geom(W, H, model="HitNet") as g:
            g.set_calibration(c)
            g.set_option("exact_div", exact)
            g.rectify_preprocess(img[None], img[None], rectify=False)
            g.set_disparity(d)
            pts, _ = g.reproject(1, DENSE)
            want, _ = O.reproject_dense(O.HITNET, d[0, 0], img, c["Q"])
            assert_close_points(pts[0], want, f"special values exact_div={exact}")
            cp, cc = g.reproject(1, COMPACT)
            wc, k, _ = O.reproject_compact(O.HITNET, d[0, 0], img, c["Q"])
            assert cc[0] == k
            assert_close_points(cp[0, :k], wc, "special values compact")


def test_reproject_tiny_disparities_with_large_Q(demo_pair):
    """Quotients that overflow: with Q entries near the 2^20 limit of the shared-reciprocal path the numerators reach 2^30, and
    a disparity around 2^-99 makes n/w exceed FLT_MAX.  IEEE division answers +-inf there; the fast path must too (its
    correction step would otherwise evaluate inf - inf = NaN), and on both sides of the window where it hands over to the
    exact division (|w| = 2^-60, 2^60) results must stay within the gate."""
    from stereoalgorithms_b200.geometry import DENSE
    W, H = 640, 480
    Q = np.array([[1.0e6, 3.0e5, 0, -2.9e5], [-2.0e5, 9.9e5, 0, 1.0e6], [1.0e3, -2.0e3, 0, 5.4e5], [0, 0, 1.0, 0]])
    vals = np.array([2.0 ** -99, -2.0 ** -99, 1.7e-30, 2.0 ** -98, 1e-31, 1e-29, 1e-28, 2.0 ** -100, 2.0 ** -101, 2.0 ** -60, 2.0 ** -61, 1.5 * 2.0 ** -60,
                     2.0 ** 60, 2.0 ** 61, 0.75 * 2.0 ** 60, 1e-20, 1e20, 3.0, 2.0 ** -126, 1e-38], np.float32)
    d = np.resize(vals, H * W).reshape(H, W).astype(np.float32)
    img = demo_pair[0]
    for exact in (0, 1):
        with geom(W, H) as g:
            g.set_Q(Q)
            g.set_option("exact_div", exact)
            g.rectify_preprocess(img[None], img[None], rectify=False)
            g.set_disparity(d[None, None])
            pts, _ = g.reproject(1, DENSE)
            want, _ = O.reproject_dense(O.FASTACVNET_PLUS, d, img, Q)
            assert np.isinf(want[:, :3]).any() and np.isfinite(want[:, :3]).any()
            assert_close_points(pts[0], want, f"tiny disparities, exact_div={exact}")


def test_side_by_side_capture_frame(calibs, demo_pair):
    """the 1280x480 side-by-side capture, split by pitch/offset instead of two clone()s."""
    import torch
    c = calibs["calib_b"]
    l, r = demo_pair
    sbs = np.ascontiguousarray(np.concatenate([l, r], axis=1))[None]          # [1, 480, 1280, 3]
    sbs = np.concatenate([sbs, sbs[:, ::-1]], 0)                               # second frame: upside down
    with geom(640, 480, model="FastACVNet_plus", max_batch=2) as g:
        g.set_calibration(c)
        for frames in (sbs, torch.from_numpy(sbs).cuda()):                     # host and device resident
            g.rectify_preprocess_side_by_side(frames)
            for i in range(2):
                li, ri = sbs[i, :, :640], sbs[i, :, 640:]
                rl, rr = O.rectify_pair(c, np.ascontiguousarray(li), np.ascontiguousarray(ri))
                assert np.array_equal(g.rectified(2, 0)[i], rl) and np.array_equal(g.rectified(2, 1)[i], rr)


@pytest.mark.parametrize("size", [(1242, 375), (638, 50), (333, 77), (67, 19), (131, 33), (1241, 40)])
@pytest.mark.parametrize("l2_prefetch", [2, 0])
@pytest.mark.parametrize("source", ["host", "host_batch_kernel", "device_dense", "device_pitched"])
def test_rectified_unaligned_widths_batches(calibs, size, l2_prefetch, source):
    """Widths that are not a multiple of 4: rows (and, for odd W*H, frames) start at every byte alignment, so the batch
    kernel's funnel-shifted loads and its run-wise aligned-word stores (head/tail bytes at warp and row ends, partial tiles
    at the right edge) see all four cases; several frames per CTA; tensors of the last frame checked too.
    source: host frames are uploaded with rows padded to 16 bytes, so with bulk = 2 even widths take the tensor-copy kernel
    ("host"; with bulk = 0 the batch kernel: "host_batch_kernel"); packed device frames take the batch kernel ("device_dense");
    device frames with a 16-byte-aligned row pitch (cudaMallocPitch style) and bulk = 2 the tensor-copy kernel again
    ("device_pitched").  (By default widths that are not a multiple of 16 stay on the batch kernel: measured faster.)"""
    W, H = size
    s = W / 640
    c = O.scale_calibration(calibs["calib_a"], s, crop_y=(480 * s - H) / 2)
    n = 6
    rng = np.random.default_rng(W * 1000 + H)
    L = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    R = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    (l1, l2), (r1, r2) = O.rectify_maps(c, (W, H))
    with geom(W, H, model="HitNet", max_batch=n) as g:
        g.set_option("l2_prefetch", l2_prefetch)
        g.set_option("batch_ctas", 1)                     # all frames in one CTA
        g.set_option("bulk", {"host": 2, "host_batch_kernel": 0, "device_dense": 1, "device_pitched": 2}[source])   # 2: tensor-copy kernel for any even width
        g.set_calibration(c)
        if source.startswith("host"):
            g.rectify_preprocess(L, R)
        elif source == "device_dense":
            import torch
            g.rectify_preprocess(torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda())
        else:
            import torch
            pitch = (3 * W + 15) // 16 * 16 + 16
            dl = torch.full((n, H, pitch), 0x5A, dtype=torch.uint8, device="cuda")      # padding bytes that must never be sampled
            dr = torch.full((n, H, pitch), 0xA5, dtype=torch.uint8, device="cuda")
            dl[:, :, :3 * W] = torch.from_numpy(L.reshape(n, H, 3 * W)).cuda()
            dr[:, :, :3 * W] = torch.from_numpy(R.reshape(n, H, 3 * W)).cuda()
            g.rectify_preprocess(dl, dr, n=n, pitch=pitch)
        gl, gr = g.rectified(n, 0), g.rectified(n, 1)
        for i in range(n):
            assert np.array_equal(gl[i], OC.remap(L[i], l1, l2)), (size, i)
            assert np.array_equal(gr[i], OC.remap(R[i], r1, r2)), (size, i)
        t = g.tensor(n, 0)[n - 1]                         # HitNet: 6 planes, left RGB then right RGB
        want = O.preprocess(O.HITNET, OC.remap(L[n - 1], l1, l2), OC.remap(R[n - 1], r1, r2), np.float32)
        assert np.array_equal(t, want)


@pytest.mark.parametrize("taper", [50, 95])
@pytest.mark.parametrize("size,ctas", [((640, 480), 4), ((640, 480), 0), ((1242, 375), 8)])
def test_rectified_tapered_chunks(calibs, size, ctas, taper):
    """chunk_taper: frame chunks of unequal length (long ones launched first); every frame must still be produced exactly once."""
    W, H = size
    s = W / 640
    c = O.scale_calibration(calibs["calib_a"], s, crop_y=(480 * s - H) / 2) if size != (640, 480) else calibs["calib_a"]
    n = 41
    rng = np.random.default_rng(taper + W)
    base = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    L = np.stack([np.roll(base, 5 * i, axis=1) for i in range(n)])
    (l1, l2), (r1, r2) = O.rectify_maps(c, (W, H))
    with geom(W, H, model="RAFTStereo", max_batch=n) as g:
        g.set_option("batch_ctas", ctas)
        g.set_option("chunk_taper", taper)
        g.set_calibration(c)
        g.rectify_preprocess(L, L)
        gl, gr = g.rectified(n, 0), g.rectified(n, 1)
        for i in range(n):
            assert np.array_equal(gl[i], OC.remap(L[i], l1, l2)), i
            assert np.array_equal(gr[i], OC.remap(L[i], r1, r2)), i


@pytest.mark.parametrize("ctas_per_sm", [1, 16, 1000])
@pytest.mark.parametrize("bulk", [1, 0, -1])
def test_rectified_long_batches_all_chunkings(calibs, demo_pair, ctas_per_sm, bulk):
    """The batch-loop kernel with many frames per CTA (register prefetch of f+1, L2 prefetch of f+2, both shared-memory
    buffers in rotation), with the default chunking, and with one frame per CTA; tensors checked too."""
    c = calibs["calib_c_rational"]                       # strong rotation: large footprints, some tiles unstaged
    n = 37
    rng = np.random.default_rng(5)
    L = np.stack([np.roll(demo_pair[0], int(rng.integers(0, 640)), axis=1) for _ in range(n)])
    R = np.stack([np.roll(demo_pair[1], int(rng.integers(0, 480)), axis=0) for _ in range(n)])
    (l1, l2), (r1, r2) = O.rectify_maps(c, (640, 480))
    lut = O.norm_lut(O.FASTACVNET_PLUS)
    with geom(640, 480, model="FastACVNet_plus", max_batch=n) as g:
        g.set_option("batch_ctas", ctas_per_sm)
        g.set_option("bulk", max(bulk, 0))
        g.set_option("no_fixed_size", 1 if bulk < 0 else 0)        # -1: the batch kernel's run-time-size instantiation
        g.set_calibration(c)
        g.rectify_preprocess(L, R)
        gl, gr, tl = g.rectified(n, 0), g.rectified(n, 1), g.tensor(n, 0)
        for i in range(n):
            wl = OC.remap(L[i], l1, l2)
            assert np.array_equal(gl[i], wl), i
            assert np.array_equal(gr[i], OC.remap(R[i], r1, r2)), i
            if i % 6 == 0:
                assert np.array_equal(tl[i], np.stack([lut[ch][wl[..., 2 - ch]] for ch in range(3)]))


@pytest.mark.parametrize("n_handles,n", [(2, 9), (3, 10), (4, 3)])
def test_run_batch_multi_equals_one_handle_bit_for_bit(calibs, demo_pair, n_handles, n):
    """sa_run_batch_multi cuts a host batch frame-wise over several handles (one per GPU, one host thread each) and must give
    exactly what sa_run_batch gives for the whole batch on one handle - point clouds, disparity (the stub numbers frames across
    shards), rectified images, heat maps.  Handles go to distinct GPUs when the box has them, else they share GPU 0 (the
    partition, the pointer arithmetic and the threading are the same).  (4, 3): more handles than pairs, one shard is empty."""
    import torch
    from stereoalgorithms_b200.geometry import PinnedArray, run_batch_multi
    c = calibs["calib_b"]
    W, H = 640, 480
    rng = np.random.default_rng(77)
    L = PinnedArray((n, H, W, 3), np.uint8)
    R = PinnedArray((n, H, W, 3), np.uint8)
    for i in range(n):
        L.array[i] = np.roll(demo_pair[0], int(rng.integers(0, W)), axis=1)
        R.array[i] = np.roll(demo_pair[1], int(rng.integers(0, W)), axis=1)
    ngpu = torch.cuda.device_count()
    per = -(-n // n_handles)

    def outputs():
        return dict(pointcloud=np.zeros((n, W * H, 6), np.float32), disparity=np.zeros((n, H, W), np.float32),
                    rect_left=np.zeros((n, H, W, 3), np.uint8), rect_right=np.zeros((n, H, W, 3), np.uint8), heatmap=np.zeros((n, H, W, 3), np.uint8))

    one, many = outputs(), outputs()
    with geom(W, H, model="RAFTStereo", max_batch=n) as g:
        g.set_calibration(c)
        g.set_option("chunk", 4)
        g.run_batch(L.array, R.array, n, rectify=True, stub_seed=99, **one)
    gs = [geom(W, H, model="RAFTStereo", max_batch=per, gpu_id=i % ngpu) for i in range(n_handles)]
    try:
        for g in gs:
            g.set_calibration(c)
            g.set_option("chunk", 2)
        for _ in range(2):                                    # twice: nothing of the first call may linger
            run_batch_multi(gs, L.array, R.array, n, rectify=True, stub_seed=99, **many)
        # too many pairs for the handles: a clear error, nothing crashes
        from stereoalgorithms_b200.geometry import StereoGeomError
        if n > n_handles:
            with pytest.raises(StereoGeomError, match="max_batch"):
                run_batch_multi(gs[:1], L.array, R.array, n, rectify=True, stub_seed=99, **outputs())
    finally:
        for g in gs:
            g.close()
    assert (one["disparity"] > 0).mean() > 0.9
    for k in one:
        assert np.array_equal(one[k], many[k], equal_nan=True), k
    (l1, l2), _ = O.rectify_maps(c, (W, H))
    assert np.array_equal(many["rect_left"][n - 1], OC.remap(L.array[n - 1], l1, l2))
    want, _ = O.reproject_dense(O.FASTACVNET_PLUS, many["disparity"][n - 1], many["rect_left"][n - 1], c["Q"])
    assert_close_points(many["pointcloud"][n - 1], want, "last frame of the last shard")


def test_engine_callback_sits_between_the_two_halves(calibs, demo_pair):
    """sa_set_engine: the slot of context->enqueueV3(stream) (TRTFastACVNet_plus.cpp:152).  The callback gets the device
    tensors of its chunk and the disparity slot, on the library's stream; here it checks the tensors it is handed and
    copies a prepared 'network output' into place with a stream-ordered device-to-device copy."""
    import torch
    from cuda.bindings import runtime as rt
    c = calibs["calib_a"]
    n, W, H = 7, 640, 480
    L = np.stack([np.roll(demo_pair[0], 11 * i, axis=1) for i in range(n)])
    R = np.stack([np.roll(demo_pair[1], 11 * i, axis=1) for i in range(n)])
    D = np.stack([O.synthetic_disparity(W, H, 300 + i) for i in range(n)])
    net_out = torch.from_numpy(D).cuda()
    calls = []

    def engine(stream, first_frame, m, left_tensor, right_tensor, disparity):
        calls.append((first_frame, m, left_tensor, right_tensor))
        src = net_out.data_ptr() + first_frame * W * H * 4
        (err,) = rt.cudaMemcpyAsync(disparity, src, m * W * H * 4, rt.cudaMemcpyKind.cudaMemcpyDeviceToDevice, stream)
        return 0 if err == rt.cudaError_t.cudaSuccess else 1

    pts = np.zeros((n, W * H, 6), np.float32)
    disp = np.zeros((n, H, W), np.float32)
    with geom(W, H, model="FastACVNet_plus", max_batch=8) as g:
        g.set_calibration(c)
        g.set_option("chunk", 3)
        g.set_engine(engine)
        g.run_batch(L, R, n, rectify=True, stub_seed=0, pointcloud=pts, disparity=disp)
        assert [(f, m) for f, m, _, _ in calls] == [(0, 3), (3, 3), (6, 1)]
        t0 = g.input_tensor_ptr(0)
        assert calls[1][2] == t0 + 3 * 3 * W * H * 4 and calls[1][3] == g.input_tensor_ptr(1) + 3 * 3 * W * H * 4
        assert np.array_equal(disp, D)
        for i in (0, 4, 6):
            ref = O.run_pair(O.FASTACVNET_PLUS, c, L[i], R[i], D[i])
            assert_close_points(pts[i], ref["points"], f"engine frame {i}")
        g.set_engine(None)
