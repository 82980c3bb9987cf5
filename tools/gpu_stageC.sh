python - <<'PY'
import numpy as np, torch, time, sys
sys.path.insert(0,'.')
from stereoalgorithms_b200.geometry import StereoGeometry, DENSE, COMPACT
from stereoalgorithms_b200 import _lib as L
for (W,H,n) in [(640,480,64),(1920,1080,32)]:
    g=StereoGeometry(W,H,model="FastACVNet_plus",max_batch=n,flags=L.FLAG_TIMING)
    g.set_Q(np.array([[1,0,0,-W/2],[0,1,0,-H/2],[0,0,0,500.],[0,0,0.0166,0]]))
    img=torch.randint(0,256,(n,H,W,3),dtype=torch.uint8,device='cuda')
    g.rectify_preprocess(img,img,rectify=False)
    g.stub_disparity(n,3)
    out=torch.empty((n,H*W,6),dtype=torch.float32,device='cuda'); cnt=torch.empty(n,dtype=torch.int32,device='cuda'); hm=torch.empty((n,H,W,3),dtype=torch.uint8,device='cuda')
    for mode,name in ((DENSE,'dense'),(COMPACT,'compact')):
        for _ in range(5): g.reproject(n,mode,out=out,counts=cnt,to_host=False); g.heatmap(n,out=hm,to_host=False)
        g.synchronize(); g.reset_timing()
        for _ in range(20): g.reproject(n,mode,out=out,counts=cnt,to_host=False); g.heatmap(n,out=hm,to_host=False)
        rp,c1=g.stage_time(L.STAGE_REPROJECT); ht,c2=g.stage_time(L.STAGE_HEATMAP)
        npx=W*H*n
        print(f"{W}x{H} n={n} {name}: reproject {rp/c1:.4f} ms ({31*npx/(rp/c1*1e-3)/1e9:.0f} GB/s dense-equivalent), heatmap+fold {ht/c2:.4f} ms ({7*npx/(ht/c2*1e-3)/1e9:.0f} GB/s), valid frac {cnt.float().mean().item()/(W*H):.3f}")
    g.close()
PY
