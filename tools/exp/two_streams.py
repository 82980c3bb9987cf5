"""Experiment: stage A of one half-batch concurrent with stage B of the other (two handles, two streams) vs one 64-pair batch."""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")  # run from the repo root
from oracle import oracle as O
from stereoalgorithms_b200.geometry import StereoGeometry, DENSE, F16

c = O.load_calibration_yaml("tests/golden/calib_a.yml")
W, H = 640, 480
def make(n):
    g = StereoGeometry(W, H, model="RAFTStereo", dtype=F16, max_batch=n)
    g.set_calibration(c); g.stub_disparity(n, 5); g.synchronize()
    img = torch.randint(0, 256, (n, H, W, 3), dtype=torch.uint8, device="cuda")
    return g, img
def timeit(fn, steps=200):
    for _ in range(20): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    for _ in range(steps): fn()
    e1.record(); torch.cuda.synchronize()
    return (time.perf_counter() - t0) / steps * 1e3
g, img = make(64)
s0 = torch.cuda.Stream()
def one():
    g.rectify_preprocess(img, img, n=64, stream=s0.cuda_stream); g.reproject(64, DENSE, to_host=False, stream=s0.cuda_stream)
print("one stream, 64 pairs: %.4f ms/step" % timeit(one))
for parts in (2, 4):
    hs = [make(64 // parts) for _ in range(parts)]
    ss = [torch.cuda.Stream() for _ in range(parts)]
    def many():
        for (gg, im), s in zip(hs, ss):
            gg.rectify_preprocess(im, im, n=64 // parts, stream=s.cuda_stream)
        for (gg, im), s in zip(hs, ss):
            gg.reproject(64 // parts, DENSE, to_host=False, stream=s.cuda_stream)
    def many2():
        for (gg, im), s in zip(hs, ss):
            gg.rectify_preprocess(im, im, n=64 // parts, stream=s.cuda_stream)
            gg.reproject(64 // parts, DENSE, to_host=False, stream=s.cuda_stream)
    print("%d streams x %d pairs, A A.. B B..: %.4f ms/step;  A B A B..: %.4f ms/step" % (parts, 64 // parts, timeit(many), timeit(many2)))
