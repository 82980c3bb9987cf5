# 1242x375 (rows 2-byte aligned) tuning sweeps: bash tools/gpu_kitti.sh
run() { python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e --workload ${WL:-fastacv_kitti_1242x375_b256_f32} "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('${WL:-kitti} $*', round(d['value']), {k:(round(v['ms_per_launch'],4), round(v['GBps'])) for k,v in d['roofline']['stages'].items()})"; }
for pf in 0 2 3; do run --opt l2_prefetch=$pf; done
for c in 32 64 256; do run --opt batch_ctas=$c; done
