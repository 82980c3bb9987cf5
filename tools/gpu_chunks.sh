# wave-quantisation sweep of the frame chunking on the headline workload
for c in 16 20 24 28 32 36 44 52; do
python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-e2e --workload ${WL:-raft_realtime_640x480_b64_f16} --opt batch_ctas=$c 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('batch_ctas=$c', round(d['value']), {k:(round(v['ms_per_launch'],4), round(v['GBps'])) for k,v in d['roofline']['stages'].items()})"
done
