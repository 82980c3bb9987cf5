#!/usr/bin/env python
"""Batch-1 device-resident step (stage A -> stub -> stage B [-> stage C]) issued call by call against the same step captured once
in a CUDA graph and replayed (tests/test_gpu_compat.py::test_device_resident_step_can_be_captured_...): wall clock per step."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from oracle import oracle as O
    from stereoalgorithms_b200.geometry import DENSE, F16, StereoGeometry
    out = {}
    for n in (1, 8):
        W, H = 640, 480
        dL = torch.randint(0, 256, (n, H, W, 3), dtype=torch.uint8, device="cuda")
        dR = torch.randint(0, 256, (n, H, W, 3), dtype=torch.uint8, device="cuda")
        pts = torch.zeros((n, W * H, 6), dtype=torch.float32, device="cuda")
        with StereoGeometry(W, H, model="FastACVNet_plus", max_batch=n) as g:
            g.set_calibration(O.load_calibration_yaml(os.path.join(ROOT, "tests", "golden", "calib_a.yml")))
            s = torch.cuda.Stream()
            sp = s.cuda_stream

            def step():
                g.rectify_preprocess(dL, dR, n=n, stream=sp)
                g.stub_disparity(n, seed=11, stream=sp)
                g.reproject(n, DENSE, out=pts, to_host=False, stream=sp)

            with torch.cuda.stream(s):
                for _ in range(20):
                    step()
            s.synchronize()
            iters = 2000
            t0 = time.perf_counter()
            for _ in range(iters):
                step()
            s.synchronize()
            eager = (time.perf_counter() - t0) / iters
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=s):
                step()
            for _ in range(20):
                graph.replay()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(iters):
                graph.replay()
            torch.cuda.synchronize()
            replay = (time.perf_counter() - t0) / iters
            out[f"n={n}"] = {"calls_us_per_step": eager * 1e6, "graph_replay_us_per_step": replay * 1e6}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
