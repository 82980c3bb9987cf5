import hashlib
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def md5(a) -> str:
    return hashlib.md5(np.ascontiguousarray(a).tobytes()).hexdigest()


def read_png(path):
    """Decode a golden PNG.  cv2 is in the image; the fallback is a minimal zlib PNG reader so the
    fixtures do not depend on it."""
    try:
        import cv2
        return cv2.imread(path, cv2.IMREAD_COLOR)
    except ImportError:                                      # pragma: no cover
        raise pytest.skip.Exception("cv2 needed to decode golden PNGs")


@pytest.fixture(scope="session")
def golden():
    with open(os.path.join(GOLDEN, "golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def demo_pair(golden):
    l = read_png(os.path.join(GOLDEN, "left0.png"))
    r = read_png(os.path.join(GOLDEN, "right0.png"))
    assert md5(l) == golden["left0_md5"] and md5(r) == golden["right0_md5"]
    return l, r


@pytest.fixture(scope="session")
def calibs():
    from oracle import oracle as O
    return {n: O.load_calibration_yaml(os.path.join(GOLDEN, n + ".yml"))
            for n in ("calib_a", "calib_b", "calib_c_rational")}


def scaled_case(calibs, name):
    """The scaled calibrations of oracle/gen_golden.py."""
    from oracle import oracle as O
    spec = {
        "kitti_1242x375": ("calib_a", 1242 / 640, (1242, 375)),
        "hd_1920x1080": ("calib_a", 3.0, (1920, 1080)),
        "hd_c_1920x1080": ("calib_c_rational", 3.0, (1920, 1080)),
        "uhd_3840x2160": ("calib_a", 6.0, (3840, 2160)),
    }[name]
    base, s, (W, H) = spec
    return O.scale_calibration(calibs[base], s, crop_y=(480 * s - H) / 2), (W, H)
