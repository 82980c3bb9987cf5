import numpy as np, sys
sys.path.insert(0, ".")
from oracle import oracle as O
from stereoalgorithms_b200.geometry import StereoGeometry
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1
calib = sys.argv[2] if len(sys.argv) > 2 else "tests/golden/calib_a.yml"
c = O.load_calibration_yaml(calib)
W, H = 640, 480
imgs = np.stack([O.synthetic_image(W, H, 7 + i) for i in range(n)])
g = StereoGeometry(W, H, max_batch=n, model="RAFTStereo")
g.set_option("bulk", 1)
g.set_calibration(c)
g.rectify_preprocess(imgs, imgs[::-1].copy()); g.synchronize()
for eye, (k, d, r, p) in enumerate([("intrinsic_left", "distCoeffs_left", "R_L", "P1"), ("intrinsic_right", "distCoeffs_right", "R_R", "P2")]):
    got = g.rectified(n, eye)
    m1, m2 = O.init_undistort_rectify_map(c[k], c[d], c[r], c[p], (W, H))
    src = imgs if eye == 0 else imgs[::-1]
    bad = 0
    for f in range(n):
        w = O.remap_bilinear(src[f], m1, m2)
        bad += int((got[f] != w).any(-1).sum())
    print("n", n, "eye", eye, "bad px", bad)
