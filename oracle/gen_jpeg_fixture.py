#!/usr/bin/env python
"""Makes tests/golden/capture_1280x480_{420,444,prog}.jpg and capture_1280x480.npz (test infrastructure; needs cv2).

The reference's camera delivers MJPG frames, 1280x480 with the two eyes side by side (usb_test.py:7-11), and cv2 (libjpeg-turbo)
decodes them on the CPU.  The fixtures are a synthetic frame of that shape built from the reference's demo pair (tests/golden/
left0.png | right0.png, 640x480 each), encoded by cv2 as baseline 4:2:0 (what a UVC camera sends), baseline 4:4:4 and progressive
4:2:0; the .npz holds, per file, cv2's decode reduced to 8x8 block means (60 x 160 x 3 float32) - small enough to commit, and
what the GPU test compares nvJPEG's decode with (two decoders agree to a fraction of a grey level per block, not per pixel).

    python oracle/gen_jpeg_fixture.py
"""
import os

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "tests", "golden")


def block_means(img: np.ndarray) -> np.ndarray:
    h, w, c = img.shape
    return img.reshape(h // 8, 8, w // 8, 8, c).astype(np.float32).mean(axis=(1, 3))


def main():
    left = cv2.imread(os.path.join(G, "left0.png"), cv2.IMREAD_COLOR)
    right = cv2.imread(os.path.join(G, "right0.png"), cv2.IMREAD_COLOR)
    frame = np.concatenate([left, right], axis=1)
    assert frame.shape == (480, 1280, 3), frame.shape
    variants = {
        "420": [cv2.IMWRITE_JPEG_QUALITY, 85],
        "444": [cv2.IMWRITE_JPEG_QUALITY, 92, cv2.IMWRITE_JPEG_SAMPLING_FACTOR, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_444],
        "prog": [cv2.IMWRITE_JPEG_QUALITY, 80, cv2.IMWRITE_JPEG_PROGRESSIVE, 1],
    }
    out = {}
    for tag, params in variants.items():
        ok, buf = cv2.imencode(".jpg", frame, params)
        assert ok
        path = os.path.join(G, f"capture_1280x480_{tag}.jpg")
        with open(path, "wb") as f:
            f.write(buf.tobytes())
        dec = cv2.imdecode(buf, cv2.IMREAD_COLOR)
        out[tag] = block_means(dec)
        print(tag, len(buf), "bytes; decode vs source: mean abs diff", float(np.abs(dec.astype(int) - frame.astype(int)).mean()))
    np.savez_compressed(os.path.join(G, "capture_1280x480.npz"), **out)


if __name__ == "__main__":
    main()
