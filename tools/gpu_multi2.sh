# 2-GPU sanity of the torchrun paths: weak (default) and strong (--strong) scaling lines, multi-GPU tests
mkdir -p gpurun_out
N=${1:-2}
nvidia-smi -L | head -4
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 100 --warmup 10 --no-ref-kernels > gpurun_out/bench_g$N.json 2> gpurun_out/bench_g$N.err; echo "weak rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 100 --warmup 10 --strong --no-extra --no-cpu-baseline --no-ref-kernels --workload fastacv_kitti_1242x375_b256_f32 > gpurun_out/bench_strong_g$N.json 2> gpurun_out/bench_strong_g$N.err; echo "strong rc=$?"
python - <<PY
import json
for f in ("bench_g$N", "bench_strong_g$N"):
    try:
        d=json.load(open(f"gpurun_out/{f}.json")); print(f, d['n_gpus'], d['scaling'], round(d['value']), d['config'].get('workload'), 'e2e', round(d['e2e']['value']) if d.get('e2e') else None, d['e2e'].get('frac_of_host_ceiling') if d.get('e2e') else None)
    except Exception as e: print(f, 'fail', e, open(f"gpurun_out/{f}.err").read()[-800:])
PY
timeout 600 python -m pytest tests -m gpu -q --timeout 300 -k "multi" 2>&1 | tail -2
