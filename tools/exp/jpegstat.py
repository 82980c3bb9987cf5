import os, sys, numpy as np
sys.path.insert(0, '.')
from oracle import oracle as O
from stereoalgorithms_b200.geometry import StereoGeometry
G='tests/golden'
def bm(img):
    h,w,c=img.shape; return img.reshape(h//8,8,w//8,8,c).astype(np.float32).mean(axis=(1,3))
means=np.load(os.path.join(G,'capture_1280x480.npz'))
try:
    import cv2
except Exception: cv2=None
for tag in ('420','444','prog'):
    data=open(os.path.join(G,f'capture_1280x480_{tag}.jpg'),'rb').read()
    with StereoGeometry(640,480,model='RAFTStereo',max_batch=2) as g:
        g.set_calibration(O.load_calibration_yaml(os.path.join(G,'calib_a.yml')))
        g.decode_capture_jpeg([data])
        cap=g.capture(1)[0]
    d=np.abs(bm(cap)-means[tag])
    print(tag,'block means: max',d.max(),'mean',d.mean(),'per channel mean',d.mean(axis=(0,1)), 'signed', (bm(cap)-means[tag]).mean(axis=(0,1)))
    if cv2 is not None:
        ref=cv2.imdecode(np.frombuffer(data,np.uint8),cv2.IMREAD_COLOR).astype(int)
        pd=np.abs(cap.astype(int)-ref)
        print('   pixels: max',pd.max(),'mean',pd.mean(),'frac>2',(pd>2).mean())
