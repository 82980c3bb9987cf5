"""Host-side logic of the C-ABI library that needs no GPU: symbol surface, calibration parser, normalisation and
resize tables, point-cloud writer, and the "fails loudly without a device" contract.  CPU only."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import GOLDEN, ROOT
from oracle import oracle as O


@pytest.fixture(scope="module")
def lib():
    from stereoalgorithms_b200 import _lib
    return _lib.load()


def _declared_symbols():
    names = []
    for hdr in ("stereo_geom.h", "stereo_algorithms_compat.h"):
        text = open(os.path.join(ROOT, "include", hdr)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        names += re.findall(r"SA_API\s+[^;(]*?\b(sa_\w+)\s*\(", text)
    return names


def test_library_exports_every_declared_symbol(lib):
    from stereoalgorithms_b200 import _lib
    declared = _declared_symbols()
    assert len(declared) >= 40 and len(set(declared)) == len(declared)
    for name in declared:
        assert hasattr(lib, name), f"{name} is declared in include/ but not exported"
        assert name in _lib.SIGNATURES, f"{name} has no ctypes signature"
    assert set(_lib.SIGNATURES) == set(declared)
    assert lib.sa_version().startswith(b"stereo_geom_b200")


def _parse(lib, path):
    from stereoalgorithms_b200._lib import sa_calibration
    c = sa_calibration()
    err = C.create_string_buffer(512)
    rc = lib.sa_parse_calibration_yaml(path.encode(), C.byref(c), err, 512)
    return rc, c, err.value.decode()


@pytest.mark.parametrize("name", ["calib_a", "calib_b", "calib_b_alt", "calib_c_rational"])
def test_calibration_parser_matches_oracle_parser(lib, name):
    path = os.path.join(GOLDEN, name + ".yml")
    rc, c, err = _parse(lib, path)
    assert rc == 0, err
    want = O.load_calibration_yaml(path)
    pairs = {"K_l": "intrinsic_left", "K_r": "intrinsic_right", "R_L": "R_L", "R_R": "R_R", "P1": "P1", "P2": "P2", "Q": "Q", "R": "R", "T": "T"}
    for f, k in pairs.items():
        assert np.array_equal(np.array(getattr(c, f)), want[k].ravel()), (name, k)      # bit-exact doubles
    nl, nr = want["distCoeffs_left"].size, want["distCoeffs_right"].size
    assert (c.nD_l, c.nD_r) == (nl, nr) and nl in (5, 8)
    assert np.array_equal(np.array(c.D_l)[:nl], want["distCoeffs_left"].ravel()) and not np.array(c.D_l)[nl:].any()
    assert np.array_equal(np.array(c.D_r)[:nr], want["distCoeffs_right"].ravel())


def test_calibration_parser_variants_and_errors(lib, tmp_path):
    base = open(os.path.join(GOLDEN, "calib_a.yml")).read()
    # no "---", no space after '[', no trailing newline, CRLF-free numbers like "0." and "1.", reordered + unknown keys
    body = base.replace("---\n", "").replace("data: [ ", "data: [").rstrip("\n")
    body = "%YAML:1.0\nvalidROIL: [ 0, 0, 640, 480 ]\nfoo: 3\n" + body.split("\n", 1)[1] + "\nvalidROIR: [ 0, 0, 640, 480 ]"
    p = tmp_path / "variant.yml"
    p.write_text(body)
    rc, c, err = _parse(lib, str(p))
    assert rc == 0, err
    want = O.load_calibration_yaml(os.path.join(GOLDEN, "calib_a.yml"))
    assert np.array_equal(np.array(c.Q), want["Q"].ravel()) and np.array_equal(np.array(c.K_l), want["intrinsic_left"].ravel())
    # errors
    from stereoalgorithms_b200 import _lib as L
    rc, _, err = _parse(lib, str(tmp_path / "missing.yml"))
    assert rc == L.SA_ERR_IO and "cannot open" in err
    (tmp_path / "short.yml").write_text(base.replace("rows: 4", "rows: 5"))
    rc, _, err = _parse(lib, str(tmp_path / "short.yml"))
    assert rc == L.SA_ERR_PARSE and "Q" in err
    (tmp_path / "nonum.yml").write_text(base.replace("1., 0., 0.,", "1., zero, 0.,", 1))
    rc, _, err = _parse(lib, str(tmp_path / "nonum.yml"))
    assert rc == L.SA_ERR_PARSE
    (tmp_path / "noq.yml").write_text(base[:base.index("Q: !!opencv-matrix")])
    rc, _, err = _parse(lib, str(tmp_path / "noq.yml"))
    assert rc == L.SA_ERR_PARSE and "'Q'" in err
    (tmp_path / "unterminated.yml").write_text(base[:base.rindex("]")])
    rc, _, err = _parse(lib, str(tmp_path / "unterminated.yml"))
    assert rc == L.SA_ERR_PARSE


@pytest.mark.parametrize("model", [0, 1, 2, 3])
def test_norm_lut_matches_oracle(lib, model):
    lut = np.empty((3, 256), np.float32)
    assert lib.sa_norm_lut(model, lut.ctypes.data) == 0
    assert np.array_equal(lut.view(np.uint32), O.norm_lut(model).view(np.uint32))
    assert lib.sa_norm_lut(7, lut.ctypes.data) != 0


@pytest.mark.parametrize("dst,src", [(640, 1920), (480, 1080), (640, 1242), (192, 375), (1248, 1920), (640, 640), (3, 1000)])
def test_resize_table_matches_oracle(lib, dst, src):
    t = np.empty((4, dst), np.int32)
    assert lib.sa_resize_table(dst, src, t.ctypes.data) == 0
    s0, s1, c0, c1 = O._resize_axis_tab(dst, src)
    assert np.array_equal(t[0], s0) and np.array_equal(t[1], s1) and np.array_equal(t[2], c0) and np.array_equal(t[3], c1)
    assert np.all(t[2] + t[3] == 2048)


def test_pointcloud_txt_writer(lib, tmp_path):
    pts = np.array([[1.5, -0.25, 1e-7, 255, 0, 12], [np.inf, -np.inf, np.nan, 1, 2, 3], [123456.789, 1e10, 3.0, 0, 0, 0],
                    [0.1, 1 / 3, -2 / 3, 7, 8, 9]], np.float32)
    path = str(tmp_path / "pointcloud.txt")
    assert lib.sa_write_pointcloud_txt(path.encode(), pts.ctypes.data, len(pts)) == 0
    lines = open(path).read().splitlines()
    want = O.pointcloud_txt_lines(pts)
    # C++ streams print nan/inf like printf; the sign of nan is platform noise
    assert [l.replace("-nan", "nan") for l in lines] == want
    assert lib.sa_write_pointcloud_txt(str(tmp_path / "no" / "dir.txt").encode(), pts.ctypes.data, 1) != 0


def _no_gpu():
    try:
        import torch
        return not torch.cuda.is_available()
    except Exception:
        return True


@pytest.mark.skipif(not _no_gpu(), reason="only meaningful on a box without a GPU")
def test_fails_loudly_without_a_device(lib):
    from stereoalgorithms_b200 import _lib as L
    from stereoalgorithms_b200.geometry import StereoGeomError, StereoGeometry
    with pytest.raises(StereoGeomError) as e:
        StereoGeometry(640, 480)
    assert e.value.code == L.SA_ERR_CUDA and "no CPU fallback" in str(e.value)
    # the reference-named API keeps the reference's semantics: a handle comes back, Run reports -1
    from stereoalgorithms_b200.compat import FastACVNet_plus as F, Mat
    p = F.Initialize("m.onnx", 0, os.path.join(GOLDEN, "calib_b.yml"))
    assert p
    img = np.zeros((480, 640, 3), np.uint8)
    assert F.RunFastACVNet_plus_RectifyImage(p, img, img.copy(), np.empty(640 * 480 * 6, np.float32), Mat()) == -1
    assert F.Version(p) == "FastACVNet_plus_Algorithm_V1.0"
    assert F.Release(p) == 0


def test_product_does_not_import_the_oracle():
    """No file of the product package may reference oracle/ (a product path through the oracle voids parity)."""
    pkg = os.path.join(ROOT, "stereoalgorithms_b200")
    for root, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h")):
                text = open(os.path.join(root, f)).read()
                assert "oracle" not in text.lower(), f"{f} mentions the oracle"


def test_pointcloud_ply_writer(lib, tmp_path):
    pts = np.array([[1.5, -0.25, 2.0, 255, 0, 12], [np.inf, 1, 1, 1, 2, 3], [0.5, 0.25, np.nan, 9, 9, 9], [3, 4, 5, 7, 8, 9]], np.float32)
    b = str(tmp_path / "cloud.ply")
    assert lib.sa_write_pointcloud_ply(b.encode(), pts.ctypes.data, len(pts), 1, 1) == 0
    raw = open(b, "rb").read()
    head, body = raw.split(b"end_header\n", 1)
    assert b"format binary_little_endian 1.0" in head and b"element vertex 2\n" in head and b"property uchar blue" in head
    rec = np.frombuffer(body, dtype=np.dtype([("xyz", "<f4", 3), ("rgb", "u1", 3)]))
    assert len(rec) == 2
    assert np.array_equal(rec["xyz"], pts[[0, 3], :3]) and np.array_equal(rec["rgb"], pts[[0, 3], 3:].astype(np.uint8))
    a = str(tmp_path / "cloud_ascii.ply")
    assert lib.sa_write_pointcloud_ply(a.encode(), pts.ctypes.data, len(pts), 0, 0) == 0
    lines = open(a).read().split("end_header\n")[1].splitlines()
    assert len(lines) == 4 and lines[0] == "1.5 -0.25 2 255 0 12" and lines[3] == "3 4 5 7 8 9"


def test_pointcloud_pcd_writer(lib, tmp_path):
    pts = np.array([[1.5, -0.25, 2.0, 255, 0, 12], [np.inf, 1, 1, 1, 2, 3], [0.5, 0.25, np.nan, 9, 9, 9], [3, 4, 5, 7, 8, 9]], np.float32)
    p = str(tmp_path / "cloud.pcd")
    assert lib.sa_write_pointcloud_pcd(p.encode(), pts.ctypes.data, len(pts), 1) == 0
    raw = open(p, "rb").read()
    head, body = raw.split(b"DATA binary\n", 1)
    assert b"VERSION 0.7" in head and b"FIELDS x y z rgb" in head and b"TYPE F F F U" in head and b"WIDTH 2\n" in head and b"POINTS 2\n" in head
    rec = np.frombuffer(body, dtype=np.dtype([("xyz", "<f4", 3), ("rgb", "<u4")]))
    assert len(rec) == 2 and np.array_equal(rec["xyz"], pts[[0, 3], :3])
    assert rec["rgb"].tolist() == [(255 << 16) | 12, (7 << 16) | (8 << 8) | 9]
    assert lib.sa_write_pointcloud_pcd(p.encode(), pts.ctypes.data, len(pts), 0) == 0
    assert b"POINTS 4\n" in open(p, "rb").read()
    assert lib.sa_write_pointcloud_pcd(None, pts.ctypes.data, 4, 0) < 0


def _lib_stereo_rectify(lib, c, size, new_size, flags, alpha):
    a = lambda x: np.ascontiguousarray(np.asarray(x, np.float64).ravel())
    K1, D1, K2, D2, R, T = (a(c[k]) for k in ("intrinsic_left", "distCoeffs_left", "intrinsic_right", "distCoeffs_right", "R", "T"))
    R1, R2, P1, P2, Q = np.zeros(9), np.zeros(9), np.zeros(12), np.zeros(12), np.zeros(16)
    r1, r2 = np.zeros(4, np.int32), np.zeros(4, np.int32)
    rc = lib.sa_stereo_rectify(K1.ctypes.data, D1.ctypes.data, len(D1), K2.ctypes.data, D2.ctypes.data, len(D2), size[0], size[1], R.ctypes.data,
                               T.ctypes.data, flags, alpha, new_size[0], new_size[1], R1.ctypes.data, R2.ctypes.data, P1.ctypes.data, P2.ctypes.data,
                               Q.ctypes.data, r1.ctypes.data, r2.ctypes.data)
    assert rc == 0
    return R1, R2, P1, P2, Q, r1, r2


def test_stereo_rectify_matches_opencv_golden(lib):
    """sa_stereo_rectify (csrc/stereo_rectify.cpp) against cv2.stereoRectify 4.13 on the 48 golden configurations,
    same gates as the oracle's own test."""
    from test_oracle_golden import _check_stereo_rectify
    assert _check_stereo_rectify(lambda c, size, new_size, flags, alpha: _lib_stereo_rectify(lib, c, size, new_size, flags, alpha)) == 48


def test_rectify_calibration_in_place_and_maps_follow(lib, tmp_path):
    """The tool's flow (Stereo_Calibration.cpp:162-179): rectify the parsed cameras for a new size, write the file, read it
    back; and bad arguments are refused."""
    from stereoalgorithms_b200._lib import sa_calibration
    rc, c, err = _parse(lib, os.path.join(GOLDEN, "calib_a.yml"))
    assert rc == 0, err
    roi_l, roi_r = np.zeros(4, np.int32), np.zeros(4, np.int32)
    assert lib.sa_rectify_calibration(C.byref(c), 640, 480, 1024, 0.0, 1280, 960, roi_l.ctypes.data, roi_r.ctypes.data) == 0
    want = O.load_calibration_yaml(os.path.join(GOLDEN, "calib_a.yml"))
    R1, R2, P1, P2, Q, w1, w2 = O.stereo_rectify(want["intrinsic_left"], want["distCoeffs_left"], want["intrinsic_right"], want["distCoeffs_right"],
                                                 (640, 480), want["R"], want["T"], 1024, 0.0, (1280, 960))
    assert np.allclose(np.array(c.P1), P1.ravel(), rtol=1e-11, atol=0) and np.allclose(np.array(c.Q), Q.ravel(), rtol=1e-11, atol=1e-15)
    assert np.allclose(np.array(c.R_L), R1.ravel(), rtol=0, atol=1e-12) and np.allclose(np.array(c.R_R), R2.ravel(), rtol=0, atol=1e-12)
    assert max(abs(int(a) - int(b)) for a, b in zip(list(roi_l) + list(roi_r), list(w1) + list(w2))) <= 1
    assert c.P1[0] > 2 * 400 and abs(c.P2[3] / c.P2[0] + 1.0 / c.Q[14]) < 1e-9                 # twice the size: focal ~2x; P2[0,3] = f * Tx, Q[3,2] = -1/Tx
    out = str(tmp_path / "rerectified.yml")
    assert lib.sa_write_calibration_yaml(out.encode(), C.byref(c)) == 0
    back = O.load_calibration_yaml(out)
    assert np.array_equal(back["P2"].ravel(), np.array(c.P2)) and np.array_equal(back["R_R"].ravel(), np.array(c.R_R))
    assert lib.sa_rectify_calibration(None, 640, 480, 0, 0.0, 0, 0, None, None) < 0
    assert lib.sa_rectify_calibration(C.byref(c), 0, 480, 0, 0.0, 0, 0, None, None) < 0


def test_netpbm_writers(lib, tmp_path):
    """Lossless stand-ins for the demos' imwrite("disparity.jpg") / imwrite("heatmap.jpg") (FastACVNet_plus/test/main.cpp:55,59)."""
    rng = np.random.default_rng(0)
    grey = rng.integers(0, 256, (7, 13), dtype=np.uint8)
    p = str(tmp_path / "d.pgm")
    assert lib.sa_write_pgm(p.encode(), grey.ctypes.data, 13, 7, 0) == 0
    raw = open(p, "rb").read()
    assert raw.startswith(b"P5\n13 7\n255\n") and raw[len(b"P5\n13 7\n255\n"):] == grey.tobytes()
    bgr = rng.integers(0, 256, (5, 9, 3), dtype=np.uint8)
    padded = np.zeros((5, 40), np.uint8)
    padded[:, :27] = bgr.reshape(5, 27)
    q = str(tmp_path / "h.ppm")
    assert lib.sa_write_ppm_bgr(q.encode(), padded.ctypes.data, 9, 5, 40) == 0       # pitched rows
    raw = open(q, "rb").read()
    assert raw.startswith(b"P6\n9 5\n255\n") and raw[len(b"P6\n9 5\n255\n"):] == bgr[..., ::-1].tobytes()
    cv2 = pytest.importorskip("cv2")                                                  # OpenCV reads both back
    assert np.array_equal(cv2.imread(p, cv2.IMREAD_UNCHANGED), grey) and np.array_equal(cv2.imread(q), bgr)
    assert lib.sa_write_pgm(p.encode(), None, 13, 7, 0) < 0 and lib.sa_write_ppm_bgr(q.encode(), bgr.ctypes.data, 9, 5, 8) < 0
    assert lib.sa_write_pgm(str(tmp_path / "no" / "dir.pgm").encode(), grey.ctypes.data, 13, 7, 0) < 0


@pytest.mark.parametrize("name", ["calib_a", "calib_c_rational"])
def test_calibration_writer_round_trip_and_scaling(lib, name, tmp_path):
    from stereoalgorithms_b200._lib import sa_calibration
    src = os.path.join(GOLDEN, name + ".yml")
    rc, c, err = _parse(lib, src)
    assert rc == 0, err
    out = str(tmp_path / "rewritten.yml")
    assert lib.sa_write_calibration_yaml(out.encode(), C.byref(c)) == 0
    # both parsers read it back bit-identically, and it is in the reference's format
    rc, c2, err = _parse(lib, out)
    assert rc == 0, err
    assert bytes(c) == bytes(c2)
    want, back = O.load_calibration_yaml(src), O.load_calibration_yaml(out)
    for k in O.CALIB_KEYS:
        assert np.array_equal(want[k].ravel(), back[k].ravel()), k
    assert open(out).read().startswith("%YAML:1.0\n---\nintrinsic_left: !!opencv-matrix\n   rows: 3\n   cols: 3\n   dt: d\n   data: [ ")
    cv2 = pytest.importorskip("cv2")                       # and cv::FileStorage itself accepts it
    fs = cv2.FileStorage(out, cv2.FILE_STORAGE_READ)
    assert np.array_equal(fs.getNode("Q").mat(), want["Q"]) and np.array_equal(fs.getNode("distCoeffs_right").mat().ravel(), want["distCoeffs_right"].ravel())
    fs.release()
    # scaling: the KITTI-size and 1080p calibrations of the benchmark configs
    for s, cy in ((1242 / 640, (480 * 1242 / 640 - 375) / 2), (3.0, 180.0)):
        sc = sa_calibration()
        assert lib.sa_scale_calibration(C.byref(c), s, 0.0, cy, C.byref(sc)) == 0
        ref = O.scale_calibration(want, s, crop_y=cy)
        for f, k in {"K_l": "intrinsic_left", "K_r": "intrinsic_right", "P1": "P1", "P2": "P2", "Q": "Q", "R_L": "R_L"}.items():
            assert np.array_equal(np.array(getattr(sc, f)), ref[k].ravel()), (s, k)
    assert lib.sa_scale_calibration(C.byref(c), -1.0, 0.0, 0.0, C.byref(sc)) != 0
