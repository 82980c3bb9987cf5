one() { python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-e2e --no-ref-kernels "$@" | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$*', round(d['value']), {k:(round(v['ms_per_launch'],4), round(v['frac'],3)) for k,v in d['roofline']['stages'].items()})"; }
one --workload exp_1248x375_b256_f32
one --workload exp_1248x384_b256_f32
one --workload exp_1248x375_b256_f32 --opt bulk=0
one --workload fastacv_kitti_1242x375_b256_f32 --pitched-rows --opt bulk=2
one --workload fastacv_kitti_1242x375_b256_f32 --pitched-rows --opt bulk=2 --opt l2_prefetch=0
one --workload exp_1240x375_b256_f32 --pitched-rows --opt bulk=2
