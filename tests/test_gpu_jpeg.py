"""SURVEY.md section 8 f3, the optional half: compressed side-by-side capture frames (the reference's camera sends MJPG 1280x480,
usb_test.py:7-11) decoded on the GPU and rectified in place.  Fixtures: oracle/gen_jpeg_fixture.py (cv2 = libjpeg-turbo).

What is checked: (1) nvJPEG's pixels against cv2's decode of the same stream, as 8x8 block means.  The two decoders are not
bit-identical: measured on these fixtures (tools/exp/jpegstat.py, B200, CUDA 12.9's nvJPEG against cv2 4.13 / libjpeg-turbo) nvJPEG's
pixels lie 0.36-0.51 grey levels BELOW libjpeg's on average (it truncates where libjpeg rounds), 0.8 % of the pixels differ by more
than 2 and single pixels by up to 10 at chroma edges of the 4:2:0 streams (4 with 4:4:4); 8x8 block means differ by 0.47 on average
and 3.1 at most.  The bounds below are those figures with head-room - a wrong channel order, a shifted eye or a broken stream are
tens of grey levels.  (2) Everything AFTER the decode against the oracle run on the very pixels the decoder produced: exact, as
everywhere else."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def geom(*a, **k):
    from stereoalgorithms_b200.geometry import StereoGeometry
    return StereoGeometry(*a, **k)


def block_means(img):
    h, w, c = img.shape
    return img.reshape(h // 8, 8, w // 8, 8, c).astype(np.float32).mean(axis=(1, 3))


@pytest.fixture(autouse=True)
def _needs_nvjpeg():
    """libnvjpeg.so.12 comes with the CUDA toolkit; on a box that has only the driver the entry points answer SA_ERR_IO
    ("JPEG decode unavailable: ...") by design - nothing to test there."""
    from stereoalgorithms_b200.geometry import StereoGeomError
    try:
        with geom(640, 480, model="RAFTStereo", max_batch=1) as g:
            g.decode_capture_jpeg([fixture("420")])
    except StereoGeomError as e:
        if "unavailable" in str(e):
            pytest.skip(str(e))
        raise


def fixture(tag):
    with open(os.path.join(GOLDEN, f"capture_1280x480_{tag}.jpg"), "rb") as f:
        return f.read()


@pytest.mark.parametrize("tag", ["420", "444", "prog"])
@pytest.mark.parametrize("n", [1, 3])
def test_jpeg_capture_frames_decode_on_the_gpu_and_rectify_in_place(calibs, tag, n):
    from stereoalgorithms_b200.geometry import LEFT, RIGHT
    want_means = np.load(os.path.join(GOLDEN, "capture_1280x480.npz"))[tag]
    c = calibs["calib_a"]
    with geom(640, 480, model="FastACVNet_plus", max_batch=4) as g:
        g.set_calibration(c)
        assert g.capture_ptr() == 0
        assert g.rectify_preprocess_jpeg([fixture(tag)] * n) == n        # n > 1: the batched decoder
        cap = g.capture(n)
        assert cap.shape == (n, 480, 1280, 3)
        for i in range(n):
            d = np.abs(block_means(cap[i]) - want_means)
            assert d.max() < 5.0 and d.mean() < 0.75, (tag, i, float(d.max()), float(d.mean()))
            assert np.array_equal(cap[i], cap[0])
        L, R = np.ascontiguousarray(cap[0][:, :640]), np.ascontiguousarray(cap[0][:, 640:])
        rl, rr = O.rectify_pair(c, L, R)
        tl, tr = O.preprocess(O.FASTACVNET_PLUS, rl, rr, np.float32)
        for i in range(n):
            assert np.array_equal(g.rectified(n, LEFT)[i], rl) and np.array_equal(g.rectified(n, RIGHT)[i], rr)
            assert np.array_equal(g.tensor(n, LEFT)[i], tl) and np.array_equal(g.tensor(n, RIGHT)[i], tr)
        # decode only, then the ordinary entry point on the capture buffer: same result
        g.decode_capture_jpeg([fixture(tag)])
        g.rectify_preprocess_side_by_side(g.capture(1), rectify=False)
        assert np.array_equal(g.rectified(1, RIGHT)[0], R)


def test_jpeg_capture_errors_are_reported(calibs):
    from stereoalgorithms_b200.geometry import StereoGeomError
    good = fixture("420")
    with geom(640, 480, model="RAFTStereo", max_batch=2) as g:
        g.set_calibration(calibs["calib_a"])
        with pytest.raises(StereoGeomError, match="JPEG frame 1"):
            g.decode_capture_jpeg([good, b"\xff\xd8 this is not a JPEG stream"])
        with pytest.raises(StereoGeomError, match="JPEG frame 0: null or empty"):
            g.decode_capture_jpeg([b"", good])
        # (a stream cut off after its header is NOT an error for nvJPEG, as for libjpeg: it decodes what is there)
        g.decode_capture_jpeg([good])                                           # the handle is still usable
    with geom(320, 240, model="RAFTStereo", max_batch=1) as g:                  # a handle of another size refuses the frame
        g.set_calibration(O.scale_calibration(calibs["calib_a"], 0.5))
        with pytest.raises(StereoGeomError, match="1280x480.*640x240"):
            g.decode_capture_jpeg([good])


def psnr(a, b):
    mse = np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)
    return 99.0 if mse == 0 else 10 * np.log10(255.0 ** 2 / mse)


def test_demo_outputs_as_jpeg_files_round_trip(tmp_path):
    """the demos store the 8-bit disparity and the heat map with cv::imwrite(".jpg") (FastACVNet_plus/test/
    main.cpp:55,59).  sa_encode_jpeg / sa_write_jpeg do that on the GPU; the stream must be a JPEG of the right size that decodes
    back (sa_decode_jpeg, and cv2 where it is installed) to the image within what quality 95 loses, host and device inputs must give
    the same bytes, and a lower quality a smaller stream."""
    import torch
    d = O.synthetic_disparity(640, 480, 3, invalid=False)
    with geom(640, 480, model="HitNet") as g:
        g.set_disparity(d[None, None])
        heat = g.heatmap(1)[0]
        grey = g.disparity_u8(1)[0]
        for name, img in (("heatmap", heat), ("disparity", grey)):
            data = g.encode_jpeg(img, 95)
            assert data[:2] == b"\xff\xd8" and data[-2:] == b"\xff\xd9", name
            assert g.encode_jpeg(torch.from_numpy(img).cuda(), 95) == data, name
            back = g.decode_jpeg(data)
            want = img if img.ndim == 3 else np.repeat(img[..., None], 3, axis=2)
            assert back.shape == want.shape, name
            assert psnr(back, want) > 33.0, (name, psnr(back, want))
            try:
                import cv2
            except ImportError:
                cv2 = None
            if cv2 is not None:
                theirs = cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_COLOR if img.ndim == 3 else cv2.IMREAD_GRAYSCALE)
                assert theirs.shape == img.shape and psnr(theirs, img) > 33.0, (name, psnr(theirs, img))
        assert len(g.encode_jpeg(heat, 50)) < len(g.encode_jpeg(heat, 95))
        path = tmp_path / "heatmap.jpg"
        g.write_jpeg(str(path), heat)
        assert path.read_bytes() == g.encode_jpeg(heat, 95)
        from stereoalgorithms_b200.geometry import StereoGeomError
        with pytest.raises(StereoGeomError, match="quality"):
            g.encode_jpeg(heat, 0)
        with pytest.raises(StereoGeomError):
            g.decode_jpeg(b"\xff\xd8 not a stream")
