# fused rectify + resize: correctness, then the resize workload fused / three launches
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 600 -k "resize or rectif or run_batch" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
for o in 0 1 "0 --opt separate_preprocess=1"; do
python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e --no-ref-kernels --workload stream_1080p_resize_640x480_b64_f32 --opt fused_resize=$o > gpurun_out/q_resize.json 2> gpurun_out/q_resize.err
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/q_resize.json")); print("fused_resize=$o", round(d['value']), d['gpu_launches'], {k:(round(v['ms_per_launch'],4), round(v['frac'],3)) for k,v in d['roofline']['stages'].items()}, d['roofline']['resize_preprocess_ms_per_step'])
except Exception as e: print('no json', e); print(open("gpurun_out/q_resize.err").read()[-1500:])
PY
done
