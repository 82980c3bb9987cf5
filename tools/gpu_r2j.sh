mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 600 -k "compact or noalloc or no_alloc or single_pass or count or bench" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python bench.py --reproject compact --no-cpu-baseline --no-extra --no-ref-kernels > gpurun_out/bench_compact.json 2> gpurun_out/bench_compact.err; echo "compact bench rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/bench_compact.json")); print(round(d['value']), {k:(round(v['ms_per_launch'],4), round(v['frac'],3)) for k,v in d['roofline']['stages'].items()}, d.get('gpu_launches'))
PY
bash tools/gpu_r2i.sh | tail -7
