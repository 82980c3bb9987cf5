"""The oracle against the golden vectors cv2 produced on the reference's fixtures
(oracle/gen_golden.py), and - where cv2 is importable - against cv2 directly.  CPU only."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, md5, scaled_case
from oracle import oracle as O
from oracle import oracle_c as OC

EYES = (("left", "intrinsic_left", "distCoeffs_left", "R_L", "P1"),
        ("right", "intrinsic_right", "distCoeffs_right", "R_R", "P2"))


def test_known_answers_from_survey(golden):
    # SURVEY.md section 8c known-answer values
    assert golden["cases"]["demo_calib_a"]["rect_left"] == "70a1e071caa25d4a5f41fb86ec8ded23"
    assert golden["cases"]["demo_calib_b"]["rect_left"] == "7866387d98d22e2208ecb572977e57ba"
    assert golden["cases"]["demo_calib_c_rational"]["rect_left"] == "342c05fa2bc8c82567b69297941d98b0"
    assert golden["jet_lut_md5"] == "ad4e9329adf93313d981a706fe6f84f6"
    assert golden["cases"]["demo_calib_a"]["oob_right"] == 165
    assert golden["cases"]["demo_calib_b"]["oob_right"] == 133


def test_jet_lut(golden):
    jet = O.jet_lut_bgr()
    assert md5(jet) == golden["jet_lut_md5"]
    assert tuple(jet[0]) == (128, 0, 0) and tuple(jet[128]) == (126, 255, 130) and tuple(jet[255]) == (0, 0, 128)


@pytest.mark.parametrize("name", ["calib_a", "calib_b", "calib_c_rational"])
@pytest.mark.parametrize("impl", ["numpy", "c"])
def test_demo_maps_and_rectified(golden, demo_pair, calibs, name, impl):
    g = golden["cases"]["demo_" + name]
    c = calibs[name]
    for (eye, k, d, r, p), img in zip(EYES, demo_pair):
        if impl == "numpy":
            m1, m2 = O.init_undistort_rectify_map(c[k], c[d], c[r], c[p], (640, 480))
            rect = O.remap_bilinear(img, m1, m2)
        else:
            m1, m2 = OC.init_maps(c[k], c[d], c[r], c[p], (640, 480))
            rect = OC.remap(img, m1, m2)
        assert md5(m1) == g["map1_" + eye] and md5(m2) == g["map2_" + eye]
        assert md5(rect) == g["rect_" + eye]
    crop = np.load(os.path.join(GOLDEN, f"crop_{name}.npz"))
    m1, m2 = O.init_undistort_rectify_map(c["intrinsic_left"], c["distCoeffs_left"], c["R_L"], c["P1"], (640, 480))
    assert np.array_equal(m1[200:232, 300:332], crop["map1_left"])
    assert np.array_equal(m2[200:232, 300:332], crop["map2_left"])


@pytest.mark.parametrize("name", ["kitti_1242x375", "hd_1920x1080", "hd_c_1920x1080", "uhd_3840x2160"])
def test_scaled_maps_and_rectified(golden, calibs, name):
    g = golden["cases"]["scaled_" + name]
    c, (W, H) = scaled_case(calibs, name)
    assert [W, H] == g["size"]
    for (eye, k, d, r, p), seed in zip(EYES, g["image_seeds"]):
        m1, m2 = OC.init_maps(c[k], c[d], c[r], c[p], (W, H))
        assert md5(m1) == g["map1_" + eye] and md5(m2) == g["map2_" + eye]
        if W <= 1920:
            n1, n2 = O.init_undistort_rectify_map(c[k], c[d], c[r], c[p], (W, H))
            assert np.array_equal(n1, m1) and np.array_equal(n2, m2)
        rect = OC.remap(O.synthetic_image(W, H, seed), m1, m2)
        assert md5(rect) == g["rect_" + eye]
    assert g["oob_left"] + g["oob_right"] > 0 or "hd_c" in name   # border rule is exercised


@pytest.mark.parametrize("tag", ["absurd_fscale_0.02", "absurd_fscale_1e-7", "absurd_k1_+50"])
def test_absurd_calibrations(golden, calibs, tag):
    g = golden["cases"][tag]
    b = calibs["calib_a"]
    K, D, R, P = b["intrinsic_left"], b["distCoeffs_left"].copy(), b["R_L"], b["P1"].copy()
    P[0, 0] *= g["fscale"]
    P[1, 1] *= g["fscale"]
    D[0, 0] += g["dk1"]
    for f in (O.init_undistort_rectify_map, OC.init_maps):
        m1, m2 = f(K, D, R, P, (640, 480))
        assert md5(m1) == g["map1"] and md5(m2) == g["map2"]
    img = O.synthetic_image(640, 480, g["image_seed"])
    assert md5(O.remap_bilinear(img, m1, m2)) == g["rect"]
    assert md5(OC.remap(img, m1, m2)) == g["rect"]


def test_resize_golden(golden):
    for tag, g in golden["resize"].items():
        img = O.synthetic_image(g["src"][0], g["src"][1], g["seed"])
        assert md5(O.resize_linear_u8(img, g["dst"])) == g["md5"], tag
        assert md5(OC.resize(img, g["dst"])) == g["md5"], tag


def _heat_inputs():
    return {"synthetic_640x480": O.synthetic_disparity(640, 480, 1234),
            "synthetic_noinvalid": O.synthetic_disparity(640, 480, 1, invalid=False),
            "negated": -O.synthetic_disparity(320, 240, 5),
            "constant": np.full((48, 64), 3.25, np.float32),
            "random": np.random.default_rng(9).normal(30, 20, (375, 1242)).astype(np.float32)}


def test_heatmap_golden(golden):
    for tag, disp in _heat_inputs().items():
        g = golden["heatmap"][tag]
        assert md5(disp) == g["disparity_md5"], tag            # the generator itself is deterministic
        assert md5(O.heatmap(disp)) == g["md5"], tag
        assert md5(OC.heatmap(disp, O.jet_lut_bgr())) == g["md5"], tag


def _check_stereo_rectify(fn):
    """fn(calib dict, size, new_size, flags, alpha) -> R1, R2, P1, P2, Q, roi1, roi2; compared with cv2.stereoRectify 4.13
    (tests/golden/stereo_rectify.json).  Gates: rotations 1e-12 absolute, P/Q 1e-11 relative to the matrix's largest
    entry, ROIs equal - except at most 2 of the 48 cases by one pixel: with alpha = 0 an inner edge lands on the image
    border by construction and the ROI is a ceil()/floor() of +-1e-13."""
    import json
    G = json.load(open(os.path.join(GOLDEN, "stereo_rectify.json")))
    calibs = {}
    off_by_one = 0
    for g in G["cases"]:
        c = calibs.setdefault(g["calib"], O.load_calibration_yaml(os.path.join(GOLDEN, g["calib"] + ".yml")))
        got = fn(c, tuple(g["size"]), tuple(g["new_size"]), g["flags"], g["alpha"])
        for i, key in enumerate(("R1", "R2", "P1", "P2", "Q")):
            ref = np.array(g[key])
            d = np.abs(np.asarray(got[i], np.float64).ravel() - ref).max()
            tol = 1e-12 if key[0] == "R" else 1e-11 * np.abs(ref).max()
            assert d <= tol, (g["calib"], g["size"], g["new_size"], g["flags"], g["alpha"], key, d)
        for key, r in (("roi1", got[5]), ("roi2", got[6])):
            r = [int(v) for v in r]
            if r != g[key]:
                assert max(abs(a - b) for a, b in zip(r, g[key])) <= 1, (g, key, r)
                off_by_one += 1
    assert off_by_one <= 2, off_by_one
    return len(G["cases"])


def test_stereo_rectify_golden():
    n = _check_stereo_rectify(lambda c, size, new_size, flags, alpha: O.stereo_rectify(
        c["intrinsic_left"], c["distCoeffs_left"], c["intrinsic_right"], c["distCoeffs_right"], size, c["R"], c["T"], flags, alpha, new_size))
    assert n == 48


def test_stereo_rectify_reproduces_the_fixture_rotations():
    """The reference's own calibration files were written by stereoRectify (Stereo_Calibration.cpp:162-176): re-deriving
    R_L / R_R from their R, T gives the stored matrices to the digits another OpenCV version agrees on."""
    for name in ("calib_a", "calib_b", "calib_c_rational"):
        c = O.load_calibration_yaml(os.path.join(GOLDEN, name + ".yml"))
        R1, R2 = O.stereo_rectify(c["intrinsic_left"], c["distCoeffs_left"], c["intrinsic_right"], c["distCoeffs_right"], (640, 480), c["R"], c["T"])[:2]
        assert np.abs(R1 - c["R_L"]).max() < 5e-6 and np.abs(R2 - c["R_R"]).max() < 5e-6, name


def _disp_u8_inputs():
    return {"special": np.load(os.path.join(GOLDEN, "disparity_u8_special.npy")),
            "synthetic_640x480": O.synthetic_disparity(640, 480, 1234),
            "negated": -O.synthetic_disparity(320, 240, 5),
            "random": np.random.default_rng(9).normal(30, 80, (375, 1242)).astype(np.float32)}


def test_disparity_u8_golden(golden):
    """imwrite("disparity.jpg", CV_32FC1) = convertTo(CV_8U): vectors produced by cv2 (oracle/gen_golden.py)."""
    for tag, disp in _disp_u8_inputs().items():
        g = golden["disparity_u8"][tag]
        got = O.disparity_u8(disp)
        assert md5(got) == g["md5"], tag
        if g["values"] is not None:
            assert got.reshape(-1).tolist() == g["values"]      # +-inf, NaN, 1e10, 2^31 -> 0; 2.5 -> 2; 255.5 -> 255


def test_norm_lut(golden):
    for m, n in O.MODEL_NAMES.items():
        assert md5(O.norm_lut(m)) == golden["norm_lut_md5"][n]
    # HitNet's double division equals the float one for all 256 values, the
    # as-written FastACV values differ from an all-f32 evaluation, and a folded FMA breaks 1e-5.
    c = np.arange(256, dtype=np.float32)
    assert np.array_equal(O.norm_lut(O.HITNET)[0], c / np.float32(255))
    f = O.norm_lut(O.FASTACVNET_PLUS)
    allf32 = ((c / np.float32(255)) - np.float32(0.485)) / np.float32(0.229)
    assert (f[0] != allf32).sum() > 100
    assert np.array_equal(O.norm_lut(O.RAFTSTEREO)[2], c)


@pytest.mark.parametrize("model", [0, 1, 2, 3])
def test_numpy_and_c_restatements_agree(calibs, demo_pair, model):
    c = calibs["calib_a"]
    rl, rr = O.rectify_pair(c, *demo_pair)
    t = O.preprocess(model, rl, rr)
    if model == O.HITNET:
        assert t.shape == (6, 480, 640)
        assert np.array_equal(t[:3], OC.preprocess_eye(model, rl)) and np.array_equal(t[3:], OC.preprocess_eye(model, rr))
    else:
        assert np.array_equal(t[0], OC.preprocess_eye(model, rl)) and np.array_equal(t[1], OC.preprocess_eye(model, rr))
    # extension a10: the f16 flavour of the C restatement rounds exactly like numpy's astype(float16)
    t16 = O.preprocess(model, rl, rr, np.float16)
    h16 = OC.preprocess_eye(model, rl, np.float16)
    assert h16.dtype == np.float16 and np.array_equal(h16, t16[:3] if model == O.HITNET else t16[0])
    d = O.synthetic_disparity(640, 480)
    if model == O.RAFTSTEREO:
        d = -d
    p, du = O.reproject_dense(model, d, rl, c["Q"])
    p2, du2 = OC.reproject(model, d, rl, c["Q"])
    assert np.array_equal(p.view(np.uint32), p2.view(np.uint32)) and np.array_equal(du, du2)
    assert np.isinf(p[:, 2]).sum() > 1000 and (p[:, 2] < 0).sum() > 100      # d==0 -> inf, d<0 -> negative z
    pc, n, _ = O.reproject_compact(model, d, rl, c["Q"])
    pc2, n2 = OC.compact(du, p)
    assert n == n2 == int(((du > 0) & np.isfinite(du)).sum()) and np.array_equal(pc.view(np.uint32), pc2.view(np.uint32))
    # z = f*B/d on valid pixels
    q = O.q_to_f32(c["Q"])
    k = O.valid_mask(du).ravel()
    assert np.allclose(pc[:, 2], q[11] / (q[14] * du.ravel()[k]), rtol=1e-6)


def test_fp16_is_rounding_of_fp32(calibs, demo_pair):
    rl, rr = O.rectify_pair(calibs["calib_b"], *demo_pair)
    a, _ = O.preprocess(O.FASTACVNET_PLUS, rl, rr, np.float16)
    b, _ = O.preprocess(O.FASTACVNET_PLUS, rl, rr, np.float32)
    assert a.dtype == np.float16 and np.array_equal(a, b.astype(np.float16))


def test_against_cv2_directly(calibs, demo_pair):
    cv2 = pytest.importorskip("cv2")
    for name, c in calibs.items():
        for (eye, k, d, r, p), img in zip(EYES, demo_pair):
            m1, m2 = cv2.initUndistortRectifyMap(c[k], c[d], c[r], c[p], (640, 480), cv2.CV_16SC2)
            o1, o2 = O.init_undistort_rectify_map(c[k], c[d], c[r], c[p], (640, 480))
            assert np.array_equal(m1, o1) and np.array_equal(m2, o2)
            assert np.array_equal(cv2.remap(img, m1, m2, cv2.INTER_LINEAR), O.remap_bilinear(img, o1, o2))


def test_yaml_parser_variants(tmp_path, calibs):
    a = calibs["calib_b"]
    alt = O.load_calibration_yaml(os.path.join(GOLDEN, "calib_b_alt.yml"))
    for k in O.CALIB_KEYS:
        assert np.array_equal(np.ravel(alt[k]), np.ravel(a[k]))
    # no space after '[', numbers like "0." / "1.", no trailing newline, extra keys, any order
    p = tmp_path / "odd.yml"
    p.write_text("%YAML:1.0\nvalidROIL: [ 0, 0, 640, 480 ]\nQ: !!opencv-matrix\n   rows: 2\n   cols: 2\n   dt: d\n"
                 "   data: [1., 0.,\n          -6.69e-02, 338.8210023]\nfoo: 3")
    q = O.load_calibration_yaml(str(p))["Q"]
    assert np.array_equal(q, np.array([[1.0, 0.0], [-6.69e-02, 338.8210023]]))


def test_pointcloud_txt_format():
    pts = np.array([[1.5, -0.25, 1e-7, 255, 0, 12], [np.inf, -np.inf, np.nan, 1, 2, 3], [123456.789, 1e10, 3.0, 0, 0, 0]], np.float32)
    lines = O.pointcloud_txt_lines(pts)
    assert lines[0] == "1.5 -0.25 1e-07 255 0 12"
    assert lines[1] == "inf -inf nan 1 2 3"
    assert lines[2] == "123457 1e+10 3 0 0 0"
