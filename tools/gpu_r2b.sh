# stage A kernel experiments: correctness of the rectify paths, then the headline / f32 / 1080p workloads
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 600 -k "rectif or tensors or batches or bench or run_batch" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
for lib in ${SA_LIBS:-default}; do
  if [ $lib != default ]; then export SA_GEOM_LIB=$PWD/tools/exp/_bin/libstereo_geom_$lib.so; fi
  for wl in ${SA_WLS:-raft_realtime_640x480_b64_f16 fastacv_640x480_b64_f32 stream_1080p_b64_f32 fastacv_kitti_1242x375_b256_f32}; do
    python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-e2e --no-ref-kernels --workload $wl > gpurun_out/q_${lib}_$wl.json 2> gpurun_out/q_${lib}_$wl.err
    python - <<PY
import json
try:
    d=json.load(open("gpurun_out/q_${lib}_$wl.json")); print("$lib $wl", round(d['value']), {k:(round(v['ms_per_launch'],4), round(v['frac'],3)) for k,v in d['roofline']['stages'].items()})
except Exception as e: print('no json', e); print(open("gpurun_out/q_${lib}_$wl.err").read()[-1500:])
PY
  done
done
