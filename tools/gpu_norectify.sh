# stage A without rectification (RunCREStereo / RunFastACVNet_plus flavour): preprocess_kernel alone
python - <<'PY'
import numpy as np, torch, sys
sys.path.insert(0,'.')
from stereoalgorithms_b200.geometry import StereoGeometry, F16, F32
from stereoalgorithms_b200 import _lib as L
for (W,H,n,model,dt) in [(640,480,64,"FastACVNet_plus",F32),(640,480,64,"RAFTStereo",F16),(1920,1080,32,"FastACVNet_plus",F32),(1242,375,128,"FastACVNet_plus",F32),(1242,375,128,"RAFTStereo",F16)]:
    g=StereoGeometry(W,H,model=model,dtype=dt,max_batch=n,flags=L.FLAG_TIMING)
    img=torch.randint(0,256,(n,H,W,3),dtype=torch.uint8,device='cuda')
    for _ in range(5): g.rectify_preprocess(img,img,rectify=False)
    g.synchronize(); g.reset_timing()
    for _ in range(20): g.rectify_preprocess(img,img,rectify=False)
    t,c=g.stage_time(L.STAGE_RECTIFY)
    tb = 2 if dt==F16 else 4
    byts = 2*n*W*H*(3+3*tb)
    print(f"{W}x{H} n={n} {model}: preprocess only {t/c:.4f} ms, {byts/(t/c*1e-3)/1e9:.0f} GB/s of {byts/1e6:.0f} MB")
    g.close()
PY
