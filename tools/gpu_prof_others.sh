# full ncu captures of the kernels outside the default bench step: preprocess4 / preprocess_lin, heat map, compaction, maps
mkdir -p gpurun_out
cat > /tmp/others.py <<'PY'
import numpy as np, torch, sys
sys.path.insert(0,'.')
from oracle import oracle as O
from stereoalgorithms_b200.geometry import StereoGeometry, DENSE, COMPACT, F32
n=64
c=O.load_calibration_yaml("tests/golden/calib_a.yml")
g=StereoGeometry(640,480,model="FastACVNet_plus",dtype=F32,max_batch=n)
g.set_calibration(c)
img=torch.randint(0,256,(n,480,640,3),dtype=torch.uint8,device='cuda')
out=torch.empty((n,640*480,6),dtype=torch.float32,device='cuda'); cnt=torch.empty(n,dtype=torch.int32,device='cuda'); hm=torch.empty((n,480,640,3),dtype=torch.uint8,device='cuda')
g.stub_disparity(n,3)
for _ in range(3):
    g.rectify_preprocess(img,img,rectify=False)
    g.reproject(n,COMPACT,out=out,counts=cnt,to_host=False)
    g.heatmap(n,out=hm,to_host=False)
g.synchronize(); g.close()
g=StereoGeometry(1242,375,model="FastACVNet_plus",dtype=F32,max_batch=n)
img=torch.randint(0,256,(n,375,1242,3),dtype=torch.uint8,device='cuda')
for _ in range(3): g.rectify_preprocess(img,img,rectify=False)
g.synchronize(); g.close()
PY
python /tmp/others.py && ncu --set full --clock-control none --import-source on -k regex:'preprocess4_kernel|preprocess_lin_kernel|heatmap_kernel|count_valid_kernel|scan_blocks_kernel|reproject_compact_kernel|reduce_minmax_kernel|init_maps_kernel|tile_info_kernel' -c 24 -o gpurun_out/others_full python /tmp/others.py > gpurun_out/others_ncu.log 2>&1
echo "capture rc=$?"; tail -3 gpurun_out/others_ncu.log
