# chunk_taper sweep: bash tools/gpu_taper.sh
for wl in raft_realtime_640x480_b64_f16 fastacv_640x480_b64_f32 fastacv_kitti_1242x375_b256_f32 stream_1080p_b64_f32; do
for t in 0 30 50 70 90; do
python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-e2e --workload $wl --opt chunk_taper=$t 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$wl taper=$t', round(d['value']), {k:(round(v['ms_per_launch'],4), round(v['GBps'])) for k,v in d['roofline']['stages'].items()})"
done; done
