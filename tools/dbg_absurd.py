import numpy as np, sys
sys.path.insert(0,".")
from oracle import oracle as O
from stereoalgorithms_b200.geometry import StereoGeometry
base=O.load_calibration_yaml("tests/golden/calib_a.yml")
img=O.synthetic_image(640,480,7)
for absurd in (0,1):
    for n in (1,2,4):
        c={k:v.copy() for k,v in base.items()}
        if absurd: c["P1"][0,0]*=0.02; c["P1"][1,1]*=0.02
        g=StereoGeometry(640,480,max_batch=n)
        g.set_option("bulk",1)
        g.set_calibration(c)
        try:
            L=np.stack([img]*n)
            g.rectify_preprocess(L,L); g.synchronize()
            r=g.rectified(n,0)[n-1]
            m1,m2=O.init_undistort_rectify_map(c["intrinsic_left"],c["distCoeffs_left"],c["R_L"],c["P1"],(640,480))
            w=O.remap_bilinear(img,m1,m2)
            print("absurd",absurd,"n",n,"ok, mismatches",int((r!=w).sum()), flush=True)
        except Exception as e:
            print("absurd",absurd,"n",n,"FAIL",str(e)[:160], flush=True); sys.exit(1)
