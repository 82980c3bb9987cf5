mkdir -p gpurun_out
for wl in ${WLS:-raft_realtime_640x480_b64_f16 fastacv_640x480_b64_f32 hitnet_640x480_b64_f32 fastacv_kitti_1242x375_b256_f32 stream_1080p_b64_f32 stream_1080p_resize_640x480_b64_f32 uhd_2160p_b16_f32}; do
  python bench.py --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 5 --workload $wl > gpurun_out/w_$wl.json 2> gpurun_out/w_$wl.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/w_$wl.json")); print("$wl", round(d['value']), 'pairs/s  step', round(d['ms_per_step'],3),'ms', {k:(round(v['ms_per_launch'],4), round(v['GBps'])) for k,v in d['roofline']['stages'].items()}, 'whole', round(d['roofline']['whole_step']['frac'],3), 'e2e', round(d['e2e']['value']), round(d['e2e']['pcie_gbs_per_gpu'],1))
except Exception as e: print('fail', e, open("gpurun_out/w_$wl.err").read()[-1500:])
PY
done
