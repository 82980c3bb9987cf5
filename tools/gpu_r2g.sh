mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 600 -k "heatmap or run_batch or compat or smoke or golden" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python bench.py --heatmap --no-cpu-baseline --no-extra --no-ref-kernels > gpurun_out/bench_heatmap.json 2> gpurun_out/bench_heatmap.err; echo "heatmap bench rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/bench_heatmap.json")); print(round(d['value']), {k:(round(v['ms_per_launch'],4), round(v['frac'],3)) for k,v in d['roofline']['stages'].items()}, d.get('gpu_launches'))
PY
