# round-end 8-GPU refresh: the driver's bench command at N = 8 (weak), the 1242x375 x256 batch strong-scaled over 8, configs[2] from one process
mkdir -p gpurun_out
N=8
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --no-ref-kernels > gpurun_out/bench_g8.json 2> gpurun_out/bench_g8.err ) 2>&1 | grep real; echo "weak rc=$?"
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus $N --steps 100 --warmup 10 --strong --no-extra --no-cpu-baseline --no-ref-kernels --workload fastacv_kitti_1242x375_b256_f32 > gpurun_out/bench_strong_g8.json 2> gpurun_out/bench_strong_g8.err ) 2>&1 | grep real; echo "strong rc=$?"
( time timeout 400 python tools/scaling_configs.py --gpus 1,8 --what kitti,stream --md gpurun_out/r02_scaling_configs_8b.md > gpurun_out/scaling8b.log 2>&1 ) 2>&1 | grep real; echo "scaling rc=$?"; tail -4 gpurun_out/scaling8b.log | cut -c1-300
python - <<PY
import json
for f in ("bench_g8", "bench_strong_g8"):
    try:
        d=json.load(open(f"gpurun_out/{f}.json")); print(f, d['n_gpus'], d['scaling'], round(d['value']), d['config'].get('workload'), 'e2e', round(d['e2e']['value']) if d.get('e2e') else None, d['e2e'].get('frac_of_host_ceiling') if d.get('e2e') else None)
    except Exception as e: print(f, 'fail', e, open(f"gpurun_out/{f}.err").read()[-800:])
PY
