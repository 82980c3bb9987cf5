#!/usr/bin/env python
"""Generate tests/golden/* and oracle/jet_lut_bgr.txt  --  TEST INFRASTRUCTURE, NOT PRODUCT.

Runs ONLY in the build container, where the reference tree (/root/reference, read-only) and
Python cv2 (OpenCV 4.13.0) are both present.  It executes the *same OpenCV functions with the
same arguments* as the reference's C++ (FastACVNet_plus/src/FastACVNet_plus_Algorithm.cpp:139-143,
FastACVNet_plus/test/main.cpp:7-24) on the reference's own fixtures and records the results as
small golden vectors, so that they can travel to the GPU box (which has no /root/reference).

    python oracle/gen_golden.py            # rewrites tests/golden/ and oracle/jet_lut_bgr.txt

Nothing produced here is reference *source*; the calibration matrices and the two decoded demo
frames are data fixtures, re-emitted through cv2 (FileStorage writer / PNG encoder).
"""
import hashlib
import json
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402

REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")

CALIBS = {  # golden name -> reference fixture
    "calib_a": "Stereo_Calibration/StereoCalibration.yml",
    "calib_b": "FastACVNet_plus/test/StereoCalibration.yml",
    "calib_c_rational": "FastACVNet_plus/test/StereoCalibration_new.yml",
}


def md5(a: np.ndarray) -> str:
    return hashlib.md5(np.ascontiguousarray(a).tobytes()).hexdigest()


def cv_load(path):
    fs = cv2.FileStorage(path, cv2.FILE_STORAGE_READ)
    d = {k: fs.getNode(k).mat() for k in O.CALIB_KEYS}
    fs.release()
    return d


def cv_write(path, calib):
    fs = cv2.FileStorage(path, cv2.FILE_STORAGE_WRITE)
    for k in O.CALIB_KEYS:
        fs.write(k, np.asarray(calib[k], np.float64))
    fs.release()


def cv_rectify(calib, left, right):
    """RectifyImage as written (FastACVNet_plus_Algorithm.cpp:132-145)."""
    size = (left.shape[1], left.shape[0])
    mlx, mly = cv2.initUndistortRectifyMap(calib["intrinsic_left"], calib["distCoeffs_left"], calib["R_L"], calib["P1"], size, cv2.CV_16SC2)
    mrx, mry = cv2.initUndistortRectifyMap(calib["intrinsic_right"], calib["distCoeffs_right"], calib["R_R"], calib["P2"], size, cv2.CV_16SC2)
    rl = cv2.remap(left, mlx, mly, cv2.INTER_LINEAR)
    rr = cv2.remap(right, mrx, mry, cv2.INTER_LINEAR)
    return (mlx, mly), (mrx, mry), rl, rr


def cv_heatmap(disp):
    """heatmap() of the demo (FastACVNet_plus/test/main.cpp:7-24) through cv2."""
    mn, mx, _, _ = cv2.minMaxLoc(disp.reshape(disp.shape[0], -1))
    mean_mat = np.full(disp.shape, mn, np.float32)
    std_mat = np.full(disp.shape, (mx - mn) / 255, np.float32)
    norm = cv2.divide(cv2.subtract(disp, mean_mat), std_mat)
    a = cv2.convertScaleAbs(norm, alpha=1)
    return cv2.applyColorMap(a, cv2.COLORMAP_JET)


def main():
    os.makedirs(OUT, exist_ok=True)
    G = {"cv2_version": cv2.__version__, "cases": {}}

    # ---- JET LUT --------------------------------------------------------------------------
    jet = cv2.applyColorMap(np.arange(256, dtype=np.uint8).reshape(1, 256), cv2.COLORMAP_JET).reshape(256, 3)
    np.savetxt(os.path.join(HERE, "jet_lut_bgr.txt"), jet.astype(np.int64), fmt="%d")
    G["jet_lut_md5"] = md5(jet)

    # ---- fixtures: decoded demo frames + calibrations ----------------------------------------
    left = cv2.imread(f"{REF}/resource/left0.jpg")
    right = cv2.imread(f"{REF}/resource/right0.jpg")
    cv2.imwrite(os.path.join(OUT, "left0.png"), left, [cv2.IMWRITE_PNG_COMPRESSION, 9])
    cv2.imwrite(os.path.join(OUT, "right0.png"), right, [cv2.IMWRITE_PNG_COMPRESSION, 9])
    G["left0_md5"], G["right0_md5"] = md5(left), md5(right)

    calibs = {}
    for name, rel in CALIBS.items():
        c = cv_load(f"{REF}/{rel}")
        calibs[name] = c
        cv_write(os.path.join(OUT, name + ".yml"), c)              # OpenCV's own writer layout
        back = cv_load(os.path.join(OUT, name + ".yml"))
        for k in O.CALIB_KEYS:
            assert np.array_equal(back[k], c[k]), (name, k)         # %.17g-class round trip is exact
        mine = O.load_calibration_yaml(f"{REF}/{rel}")
        for k in O.CALIB_KEYS:
            assert np.array_equal(mine[k], c[k]), ("oracle yaml parser", name, k)
    # a second layout of calib_b written by the oracle's own writer (exercises another formatter)
    O.write_calibration_yaml(os.path.join(OUT, "calib_b_alt.yml"), calibs["calib_b"])
    back = O.load_calibration_yaml(os.path.join(OUT, "calib_b_alt.yml"))
    for k in O.CALIB_KEYS:
        assert np.array_equal(back[k].ravel(), np.asarray(calibs["calib_b"][k]).ravel()), k

    # ---- config 1: demo frames x 3 calibrations (maps, rectified images) ---------------------
    for name, c in calibs.items():
        (l1, l2), (r1, r2), rl, rr = cv_rectify(c, left.copy(), right.copy())
        case = dict(size=[640, 480], map1_left=md5(l1), map2_left=md5(l2), map1_right=md5(r1), map2_right=md5(r2),
                    rect_left=md5(rl), rect_right=md5(rr))
        # oracle must agree before we trust it
        (o_l, o_r) = O.rectify_maps(c, (640, 480))
        assert np.array_equal(o_l[0], l1) and np.array_equal(o_l[1], l2), name
        assert np.array_equal(o_r[0], r1) and np.array_equal(o_r[1], r2), name
        assert np.array_equal(O.remap_bilinear(left, l1, l2), rl), name
        assert np.array_equal(O.remap_bilinear(right, r1, r2), rr), name
        # border statistics (how many pixels have an out-of-image tap)
        for eye, m1 in (("left", l1), ("right", r1)):
            sx, sy = m1[..., 0].astype(int), m1[..., 1].astype(int)
            case["oob_" + eye] = int(((sx < 0) | (sx + 1 >= 640) | (sy < 0) | (sy + 1 >= 480)).sum())
        G["cases"]["demo_" + name] = case
        np.savez_compressed(os.path.join(OUT, f"crop_{name}.npz"),
                            map1_left=l1[200:232, 300:332], map2_left=l2[200:232, 300:332],
                            rect_left=rl[200:232, 300:332], rect_right=rr[0:32, 0:32])

    # ---- configs 3/4/5: scaled calibrations, synthetic images -----------------------------------
    base = calibs["calib_a"]
    scaled = {
        "kitti_1242x375": (O.scale_calibration(base, 1242 / 640, crop_y=(480 * 1242 / 640 - 375) / 2), (1242, 375)),
        "hd_1920x1080": (O.scale_calibration(base, 3.0, crop_y=(480 * 3 - 1080) / 2), (1920, 1080)),
        "hd_c_1920x1080": (O.scale_calibration(calibs["calib_c_rational"], 3.0, crop_y=(480 * 3 - 1080) / 2), (1920, 1080)),
        "uhd_3840x2160": (O.scale_calibration(base, 6.0, crop_y=(480 * 6 - 2160) / 2), (3840, 2160)),
    }
    for name, (c, (W, H)) in scaled.items():
        li, ri = O.synthetic_image(W, H, 2000), O.synthetic_image(W, H, 2001)
        (l1, l2), (r1, r2), rl, rr = cv_rectify(c, li, ri)
        (o_l, o_r) = O.rectify_maps(c, (W, H))
        nm = int((o_l[0] != l1).sum() + (o_l[1] != l2).sum() + (o_r[0] != r1).sum() + (o_r[1] != r2).sum())
        assert nm == 0, (name, nm)
        assert np.array_equal(O.remap_bilinear(li, l1, l2), rl) and np.array_equal(O.remap_bilinear(ri, r1, r2), rr), name
        case = dict(size=[W, H], map1_left=md5(l1), map2_left=md5(l2), map1_right=md5(r1), map2_right=md5(r2),
                    rect_left=md5(rl), rect_right=md5(rr), image_seeds=[2000, 2001])
        for eye, m1 in (("left", l1), ("right", r1)):
            sx, sy = m1[..., 0].astype(int), m1[..., 1].astype(int)
            case["oob_" + eye] = int(((sx < 0) | (sx + 1 >= W) | (sy < 0) | (sy + 1 >= H)).sum())
        G["cases"]["scaled_" + name] = case

    # ---- absurd calibrations: saturation / INT_MIN behaviour of the map generator --------------
    for tag, fscale, dk1 in (("absurd_fscale_0.02", 0.02, 0.0), ("absurd_fscale_1e-7", 1e-7, 0.0), ("absurd_k1_+50", 1.0, 50.0)):
        K, D, R, P = base["intrinsic_left"], base["distCoeffs_left"].copy(), base["R_L"], base["P1"].copy()
        P[0, 0] *= fscale
        P[1, 1] *= fscale
        D[0, 0] += dk1
        m1, m2 = cv2.initUndistortRectifyMap(K, D, R, P, (640, 480), cv2.CV_16SC2)
        o1, o2 = O.init_undistort_rectify_map(K, D, R, P, (640, 480))
        assert np.array_equal(o1, m1) and np.array_equal(o2, m2), tag
        img = O.synthetic_image(640, 480, 7)
        rm = cv2.remap(img, m1, m2, cv2.INTER_LINEAR)
        assert np.array_equal(O.remap_bilinear(img, m1, m2), rm), tag
        G["cases"][tag] = dict(size=[640, 480], fscale=fscale, dk1=dk1, map1=md5(m1), map2=md5(m2), rect=md5(rm), image_seed=7)

    # ---- resize (extension a8): downscales ------------------------------------------------
    G["resize"] = {}
    for tag, (W, H), (w, h), seed in (("1080p_to_640x480", (1920, 1080), (640, 480), 11),
                                        ("720p_to_640x480", (1280, 720), (640, 480), 12),
                                        ("1080p_to_1248x384", (1920, 1080), (1248, 384), 13),
                                        ("2x_1280x960_to_640x480", (1280, 960), (640, 480), 14),
                                        ("1242x375_to_640x192", (1242, 375), (640, 192), 15)):
        img = O.synthetic_image(W, H, seed)
        ref = cv2.resize(img, (w, h), interpolation=cv2.INTER_LINEAR)
        mine = O.resize_linear_u8(img, (w, h))
        assert np.array_equal(ref, mine), (tag, int((ref != mine).sum()))
        G["resize"][tag] = dict(src=[W, H], dst=[w, h], seed=seed, md5=md5(ref))
    # a smooth (non-random) image too: the rectified demo frame of calib_a upscaled source 1280x720 fixture
    big = cv2.imread(f"{REF}/resource/left.png")
    ref = cv2.resize(big, (640, 480), interpolation=cv2.INTER_LINEAR)
    assert np.array_equal(ref, O.resize_linear_u8(big, (640, 480)))

    # ---- heatmap -------------------------------------------------------------------------------
    G["heatmap"] = {}
    for tag, disp in (("synthetic_640x480", O.synthetic_disparity(640, 480, 1234)),
                      ("synthetic_noinvalid", O.synthetic_disparity(640, 480, 1, invalid=False)),
                      ("negated", -O.synthetic_disparity(320, 240, 5)),
                      ("constant", np.full((48, 64), 3.25, np.float32)),
                      ("random", np.random.default_rng(9).normal(30, 20, (375, 1242)).astype(np.float32))):
        ref = cv_heatmap(disp)
        mine = O.heatmap(disp)
        assert np.array_equal(ref, mine), (tag, int((ref != mine).sum()))
        G["heatmap"][tag] = dict(shape=list(disp.shape), md5=md5(ref), disparity_md5=md5(disp))

    # ---- disparity -> 8-bit image (what imwrite("disparity.jpg", CV_32FC1) encodes): cv2 PNG round trip ----------
    G["disparity_u8"] = {}
    special = np.array([[0.5, 1.5, 2.5, -3.2, 254.5, 255.5, 300, np.inf, -np.inf, np.nan, 0.49999997, 127.50001, 1e10, -1e10,
                         2147483648.0, 3e9, 2147483520.0, -0.5, 255.49998, 63.5]], np.float32)
    np.save(os.path.join(OUT, "disparity_u8_special.npy"), special)
    for tag, disp in (("special", special),
                      ("synthetic_640x480", O.synthetic_disparity(640, 480, 1234)),
                      ("negated", -O.synthetic_disparity(320, 240, 5)),
                      ("random", np.random.default_rng(9).normal(30, 80, (375, 1242)).astype(np.float32))):
        import tempfile
        with tempfile.TemporaryDirectory() as td:
            cv2.imwrite(os.path.join(td, "d.png"), disp)                     # same convertTo(CV_8U) as the .jpg path, lossless container
            ref = cv2.imread(os.path.join(td, "d.png"), cv2.IMREAD_UNCHANGED)
        mine = O.disparity_u8(disp)
        assert ref.dtype == np.uint8 and np.array_equal(ref.reshape(disp.shape), mine), (tag, int((ref.reshape(disp.shape) != mine).sum()))
        G["disparity_u8"][tag] = dict(shape=list(disp.shape), md5=md5(ref), values=ref.reshape(-1).tolist() if tag == "special" else None)

    # ---- stereoRectify (section 8f-2): cv2 outputs for 48 configurations ----------------------------------------
    SR = []
    worst = [0.0] * 5
    roi_mismatch = 0
    for name in ("calib_a", "calib_b", "calib_c_rational"):
        c = O.load_calibration_yaml(os.path.join(OUT, name + ".yml"))
        args = [c[k] for k in ("intrinsic_left", "distCoeffs_left", "intrinsic_right", "distCoeffs_right")]
        for size, new_size in (((640, 480), (0, 0)), ((1920, 1080), (640, 480))):
            for flags in (cv2.CALIB_ZERO_DISPARITY, 0):
                for alpha in (-1.0, 0.0, 0.5, 1.0):
                    ref = cv2.stereoRectify(*args, size, c["R"], c["T"], flags=flags, alpha=alpha, newImageSize=new_size)
                    mine = O.stereo_rectify(*args, size, c["R"], c["T"], flags, alpha, new_size)
                    for i in range(5):
                        worst[i] = max(worst[i], float(np.abs(mine[i] - ref[i]).max() / max(1.0, np.abs(ref[i]).max())))
                    roi_mismatch += tuple(mine[5]) != tuple(ref[5]) or tuple(mine[6]) != tuple(ref[6])
                    SR.append(dict(calib=name, size=list(size), new_size=list(new_size), flags=int(flags), alpha=alpha,
                                   R1=ref[0].ravel().tolist(), R2=ref[1].ravel().tolist(), P1=ref[2].ravel().tolist(), P2=ref[3].ravel().tolist(),
                                   Q=ref[4].ravel().tolist(), roi1=[int(v) for v in ref[5]], roi2=[int(v) for v in ref[6]]))
    assert max(worst) < 1e-12 and roi_mismatch <= 2, (worst, roi_mismatch)
    print("stereoRectify: oracle vs cv2, worst scaled error per output", worst, "roi mismatches", roi_mismatch, "of", len(SR))
    with open(os.path.join(OUT, "stereo_rectify.json"), "w") as f:
        json.dump(dict(opencv=cv2.__version__, cases=SR), f)

    # ---- preprocess LUTs + reproject known answers (oracle-produced; the reference has no CPU
    #      code for these stages - they are pinned on the GPU box against the reference's own
    #      kernels compiled unmodified into oracle/_ref, see oracle/build_ref.py) ------------------
    G["norm_lut_md5"] = {O.MODEL_NAMES[m]: md5(O.norm_lut(m)) for m in O.MODEL_NAMES}

    with open(os.path.join(OUT, "golden.json"), "w") as f:
        json.dump(G, f, indent=1, sort_keys=True)
    print("golden written:", OUT)
    print(json.dumps({k: v for k, v in G["cases"].items() if k.startswith("demo_")}, indent=1))


if __name__ == "__main__":
    main()
