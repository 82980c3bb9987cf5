"""Per-call latency of the reference-named entry point (batch 1, host cv::Mat-like buffers in and out), the way the
reference's demo calls it (FastACVNet_plus/test/main.cpp:42-53).  usage: python tools/compat_latency.py [iterations]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereoalgorithms_b200.compat import FastACVNet_plus as F, Mat, RAFTStereo as Rf  # noqa: E402

GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
iters = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 200
import cv2  # noqa: E402  (only to decode the demo frames)

l0, r0 = cv2.imread(os.path.join(GOLDEN, "left0.png")), cv2.imread(os.path.join(GOLDEN, "right0.png"))
for name, mod, run in (("FastACVNet_plus", F, "RunFastACVNet_plus_RectifyImage"), ("RAFTStereo", Rf, "RunRAFTStereo")):
    p = mod.Initialize("model.onnx", 0, os.path.join(GOLDEN, "calib_a.yml"))
    if "--register" in sys.argv:
        mod.SetHostRegister(p, True)       # opt-in page-locking of the reused pointcloud buffer
    pc = np.empty(640 * 480 * 6, np.float32)
    d = Mat()
    fn = getattr(mod, run)
    times = []
    for i in range(iters + 20):
        l, r = l0.copy(), r0.copy()
        t0 = time.perf_counter()
        rc = fn(p, l, r, pc, d)
        t1 = time.perf_counter()
        assert rc == 0, mod.LastError(p)
        if i >= 20:
            times.append(t1 - t0)
    times = np.array(times) * 1e3
    print(f"{name}.{run}: median {np.median(times):.3f} ms, p10 {np.percentile(times, 10):.3f}, p90 {np.percentile(times, 90):.3f} per pair "
          f"(640x480, pageable host buffers, H2D 1.8 MB + D2H 10.4 MB inside the call, synthetic-disparity stub as the network) "
          f"= {1e3 / np.median(times):.0f} pairs/s single-stream")
    mod.Release(p)
