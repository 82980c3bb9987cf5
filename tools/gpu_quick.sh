# quick correctness + bench loop: bash tools/gpu_quick.sh [pytest -k expr]
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 600 ${1:+-k "$1"} > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
for wl in raft_realtime_640x480_b64_f16 stream_1080p_b64_f32; do
python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-e2e --workload $wl > gpurun_out/q_$wl.json 2> gpurun_out/q_$wl.err; echo "bench $wl rc=$?"
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/q_$wl.json")); print({k:round(d[k],4) for k in ('value','ms_per_step')}, {k:(round(v['ms_per_launch'],4), round(v['GBps'])) for k,v in d['roofline']['stages'].items()}, d['clocks'])
except Exception as e: print('no json', e); print(open("gpurun_out/q_$wl.err").read()[-2000:])
PY
done
