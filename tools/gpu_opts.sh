# usage: bash tools/gpu_opts.sh key v1 v2 ...   : quick bench of the workloads for each value of one option
key=$1; shift
for v in "$@"; do
  for wl in raft_realtime_640x480_b64_f16 fastacv_kitti_1242x375_b256_f32 stream_1080p_b64_f32; do
    python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-e2e --workload $wl --opt $key=$v 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$key=$v', '$wl', round(d['value']), {k:(round(v['ms_per_launch'],4), round(v['GBps'])) for k,v in d['roofline']['stages'].items()})"
  done
done
