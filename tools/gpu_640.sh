# the three 640x480 workloads (the fixed-size instantiation of the batch kernel), optionally with bench options
for wl in raft_realtime_640x480_b64_f16 fastacv_640x480_b64_f32 hitnet_640x480_b64_f32; do
python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-e2e --workload $wl "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$wl $*', round(d['value']), {k:(round(v['ms_per_launch'],4), round(v['GBps'])) for k,v in d['roofline']['stages'].items()})"
done
