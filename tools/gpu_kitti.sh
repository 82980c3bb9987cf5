run() { python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e --workload ${WL:-fastacv_kitti_1242x375_b256_f32} "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('${WL:-kitti} $*', round(d['value']), {k:(round(v['ms_per_launch'],4), round(v['GBps'])) for k,v in d['roofline']['stages'].items()})"; }
for WL in fastacv_kitti_1242x375_b256_f32 exp_1240x375_b256_f32 raft_realtime_640x480_b64_f16 fastacv_640x480_b64_f32 stream_1080p_b64_f32 uhd_2160p_b16_f32; do for pf in 0 2 3; do run --opt l2_prefetch=$pf; done; done
