# persistent stage A kernel vs one work item per CTA: correctness first, then both on the four stage A workloads
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 600 -k "rectif or tensors or batches or bench or run_batch or resize" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
for opt in no_persistent_ws=0 no_persistent_ws=1; do
  for wl in raft_realtime_640x480_b64_f16 fastacv_640x480_b64_f32 stream_1080p_b64_f32 fastacv_kitti_1242x375_b256_f32; do
    python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-e2e --no-ref-kernels --workload $wl --opt $opt > gpurun_out/q_${opt}_$wl.json 2> gpurun_out/q_${opt}_$wl.err
    python - <<PY
import json
try:
    d=json.load(open("gpurun_out/q_${opt}_$wl.json")); print("$opt $wl", round(d['value']), {k:(round(v['ms_per_launch'],4), round(v['frac'],3)) for k,v in d['roofline']['stages'].items()})
except Exception as e: print('no json', e); print(open("gpurun_out/q_${opt}_$wl.err").read()[-1500:])
PY
  done
done
