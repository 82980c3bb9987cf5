#!/usr/bin/env python
"""Compressed capture frames: n MJPG frames 1280x480 -> decode on the GPU (nvJPEG) -> stage A in place,
against the reference's route (cv2.VideoCapture's CPU decode, here cv2.imdecode, then the split and the upload).

    python tools/jpeg_capture_bench.py [--n 64] [--steps 20]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=64)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--fixture", default="420")
    args = ap.parse_args()
    from oracle import oracle as O
    from stereoalgorithms_b200.geometry import StereoGeometry
    g_dir = os.path.join(ROOT, "tests", "golden")
    data = open(os.path.join(g_dir, f"capture_1280x480_{args.fixture}.jpg"), "rb").read()
    jpegs = [data] * args.n
    out = {"frames": args.n, "jpeg_bytes": len(data), "fixture": args.fixture}
    with StereoGeometry(640, 480, model="RAFTStereo", dtype=1, max_batch=args.n) as g:
        g.set_calibration(O.load_calibration_yaml(os.path.join(g_dir, "calib_a.yml")))
        for what, fn in (("decode_only", lambda: g.decode_capture_jpeg(jpegs)), ("decode_and_stage_a", lambda: g.rectify_preprocess_jpeg(jpegs))):
            for _ in range(3):
                fn()
            g.synchronize()
            t0 = time.perf_counter()
            for _ in range(args.steps):
                fn()
            g.synchronize()
            dt = (time.perf_counter() - t0) / args.steps
            out[what] = {"ms_per_batch": dt * 1e3, "frames_per_s": args.n / dt}
        try:
            import cv2
            buf = np.frombuffer(data, np.uint8)
            t0 = time.perf_counter()
            k = 0
            while time.perf_counter() - t0 < 2.0:
                img = cv2.imdecode(buf, cv2.IMREAD_COLOR)
                left, right = img[:, :640].copy(), img[:, 640:].copy()          # the reference's two clone()s
                k += 1
            dt = (time.perf_counter() - t0) / k
            out["cv2_imdecode_one_core"] = {"ms_per_frame": dt * 1e3, "frames_per_s": 1 / dt, "cv2": cv2.__version__}
        except ImportError:
            pass
    print(json.dumps(out))


if __name__ == "__main__":
    main()
