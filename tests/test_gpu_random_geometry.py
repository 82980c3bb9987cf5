"""Seeded random configurations of the whole path against the oracle: image size (multiples of 16, even, odd), batch, model, tensor
type, a perturbed calibration (extra rotation of both eyes, scaled distortion: footprints of every shape, tiles that do not fit
the staging box), kernel options.  Rectified images and tensors exact, dense points within 1e-5, compact counts and order exact.
Complements the hand-picked cases of test_gpu_parity.py: whatever the stage A dispatcher picks for a configuration has to agree."""
import numpy as np
import pytest

from oracle import oracle as O
from oracle import oracle_c as OC
from test_gpu_parity import assert_close_points, geom

pytestmark = pytest.mark.gpu

MODELS = {"RAFTStereo": O.RAFTSTEREO, "HitNet": O.HITNET, "CREStereo": O.CRESTEREO, "FastACVNet_plus": O.FASTACVNET_PLUS}


def random_case(calibs, seed):
    rng = np.random.default_rng(1000 + seed)
    kind = seed % 4
    if kind in (0, 1):
        W = 16 * int(rng.integers(3, 42) if seed % 8 else rng.integers(60, 125))   # the warp-specialised kernel's widths (every 8th: up to 1984)
    elif kind == 2:
        W = 2 * int(rng.integers(24, 330))
    else:
        W = 2 * int(rng.integers(24, 330)) + 1
    H = int(rng.integers(17, 260))
    n = int(rng.integers(1, 8))
    base = calibs[("calib_a", "calib_b", "calib_c_rational")[int(rng.integers(0, 3))]]
    s = W / 640
    c = O.scale_calibration(base, s, crop_y=(480 * s - H) / 2)
    for key in ("R_L", "R_R"):                            # up to ~3 degrees about a random axis on top of the rectifying rotation
        c[key] = O.rodrigues_to_matrix(rng.normal(0, 0.02, 3)) @ c[key]
    for key in ("distCoeffs_left", "distCoeffs_right"):
        c[key] = c[key] * rng.uniform(0.3, 1.6)
    model = list(MODELS)[int(rng.integers(0, 4))]
    dtype = ("f32", "f16")[int(rng.integers(0, 2))]
    opts = {"bulk": int(rng.integers(0, 3)), "batch_ctas": int((0, 0, 1, 4, 64)[int(rng.integers(0, 5))]), "no_fixed_size": int(rng.integers(0, 2))}
    return W, H, n, c, model, dtype, opts, rng


@pytest.mark.parametrize("seed", range(64))
def test_random_configuration_against_the_oracle(calibs, seed):
    from stereoalgorithms_b200.geometry import COMPACT, DENSE, F16, F32
    W, H, n, c, model, dtype, opts, rng = random_case(calibs, seed)
    tag = f"seed {seed}: {W}x{H} n={n} {model} {dtype} {opts}"
    L = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    R = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    (l1, l2), (r1, r2) = O.rectify_maps(c, (W, H))
    npdt = np.float16 if dtype == "f16" else np.float32
    with geom(W, H, model=model, dtype=F16 if dtype == "f16" else F32, max_batch=n) as g:
        for k, v in opts.items():
            g.set_option(k, v)
        g.set_calibration(c)
        g.rectify_preprocess(L, R)
        gl, gr = g.rectified(n, 0), g.rectified(n, 1)
        tl = g.tensor(n, 0)
        tr = g.tensor(n, 1) if model != "HitNet" else None
        for i in range(n):
            wl, wr = OC.remap(L[i], l1, l2), OC.remap(R[i], r1, r2)
            assert np.array_equal(gl[i], wl) and np.array_equal(gr[i], wr), (tag, i)
            if i in (0, n - 1):
                want = O.preprocess(MODELS[model], wl, wr, npdt)
                if model == "HitNet":
                    assert np.array_equal(tl[i], want), (tag, i)
                else:
                    assert np.array_equal(tl[i], want[0]) and np.array_equal(tr[i], want[1]), (tag, i)
        planes = 2 if model == "CREStereo" else 1
        d = np.stack([O.synthetic_disparity(W, H, 7 * seed + i) * (0.5 + i) for i in range(n)])
        net = np.zeros((n, planes, H, W), np.float32)
        net[:, 0] = -d if model == "RAFTStereo" else d            # RAFT's engine emits the negated disparity
        g.set_disparity(net)
        pts, _ = g.reproject(n, DENSE)
        i = n - 1
        wl = OC.remap(L[i], l1, l2)
        want, _ = O.reproject_dense(MODELS[model], net[i, 0], wl, c["Q"])
        assert_close_points(pts[i], want, tag + " dense")
        g.set_disparity(net)
        cpts, cnt = g.reproject(n, COMPACT)
        wantc, want_count, _ = O.reproject_compact(MODELS[model], net[i, 0], wl, c["Q"])
        assert int(cnt[i]) == want_count == len(wantc), tag
        assert_close_points(cpts[i][: len(wantc)], wantc, tag + " compact")


@pytest.mark.parametrize("seed", range(20))
def test_random_resize_configuration_against_the_oracle(calibs, seed):
    """The resize path (extension a8) on random source / network sizes: arbitrary downscale ratios (different in x and y), exact 2x
    (the rounded 2x2 mean), source widths that are multiples of 16 / even / odd, the default two launches, the one-launch fused
    kernel and round 1's three launches; rectified images, both resized images and the tensors exact."""
    from stereoalgorithms_b200.geometry import F16, F32
    rng = np.random.default_rng(5000 + seed)
    kind = seed % 4
    W = 16 * int(rng.integers(12, 70)) if kind < 2 else (2 * int(rng.integers(100, 500)) + (1 if kind == 3 else 0))
    H = int(rng.integers(60, 300))
    if seed % 5 == 0 and W % 2 == 0 and H % 2 == 0:
        dst = (W // 2, H // 2)                                   # exact 2x: OpenCV's area path
    else:
        dst = (max(8, int(W / rng.uniform(1.0, 4.0))), max(8, int(H / rng.uniform(1.0, 4.0))))
    n = int(rng.integers(1, 5))
    model = list(MODELS)[int(rng.integers(0, 4))]
    dtype = ("f32", "f16")[int(rng.integers(0, 2))]
    mode = int(rng.integers(0, 3))                               # 0 default, 1 fused, 2 separate
    s = W / 640
    c = O.scale_calibration(calibs[("calib_a", "calib_b")[int(rng.integers(0, 2))]], s, crop_y=(480 * s - H) / 2)
    tag = f"seed {seed}: {W}x{H} -> {dst[0]}x{dst[1]} n={n} {model} {dtype} mode={mode}"
    L = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    R = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    (l1, l2), (r1, r2) = O.rectify_maps(c, (W, H))
    npdt = np.float16 if dtype == "f16" else np.float32
    with geom(W, H, model=model, net_size=dst, dtype=F16 if dtype == "f16" else F32, max_batch=n) as g:
        g.set_calibration(c)
        g.set_option("fused_resize", 1 if mode == 1 else 0)
        g.set_option("separate_preprocess", 1 if mode == 2 else 0)
        g.set_option("bulk", int(rng.integers(1, 3)))
        g.rectify_preprocess(L, R)
        gl, gr = g.rectified(n, 0), g.rectified(n, 1)
        nl, nr = g.net_image(n, 0), g.net_image(n, 1)
        tl = g.tensor(n, 0)
        tr = g.tensor(n, 1) if model != "HitNet" else None
        for i in range(n):
            wl, wr = OC.remap(L[i], l1, l2), OC.remap(R[i], r1, r2)
            assert np.array_equal(gl[i], wl) and np.array_equal(gr[i], wr), (tag, i)
            sl, sr = O.resize_linear_u8(wl, dst), O.resize_linear_u8(wr, dst)
            assert np.array_equal(nl[i], sl) and np.array_equal(nr[i], sr), (tag, i)
            want = O.preprocess(MODELS[model], sl, sr, npdt)
            if model == "HitNet":
                assert np.array_equal(tl[i], want), (tag, i)
            else:
                assert np.array_equal(tl[i], want[0]) and np.array_equal(tr[i], want[1]), (tag, i)
