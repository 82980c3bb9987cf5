# any-width tensor-copy path: full GPU suite, then the 1242x375 workload pitched / dense and the others
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
one() { python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e --no-ref-kernels "$@" | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$*', round(d['value']), {k:(round(v['ms_per_launch'],4), round(v['frac'],3)) for k,v in d['roofline']['stages'].items()})"; }
one --workload fastacv_kitti_1242x375_b256_f32
one --workload fastacv_kitti_1242x375_b256_f32 --dense-rows
one --workload raft_realtime_640x480_b64_f16
one --workload stream_1080p_b64_f32
