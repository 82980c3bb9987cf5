# what the 1242-wide workload pays for: 1240 (A4 loads, sector-aligned tensor rows), 1244 (A4 loads, misaligned tensor rows), 1242 (unaligned loads, misaligned rows),
# 1242 with f16 tensors (half the tensor bytes, rows 4-byte-misaligned), all dense and on the batch kernel (bulk=0 forces it for 1248)
mkdir -p gpurun_out
for wl in exp_1240x375_b256_f32 exp_1244x375_b256_f32 fastacv_kitti_1242x375_b256_f32 exp_1242x375_b256_f16 exp_1248x375_b256_f32; do
  for opt in bulk=0; do
    python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e --no-ref-kernels --workload $wl --opt $opt > gpurun_out/k3_$wl.json 2> gpurun_out/k3_$wl.err
    python - <<PY
import json
try:
    d=json.load(open("gpurun_out/k3_$wl.json")); print("$wl $opt", {k:(round(v['ms_per_launch'],4), round(v['frac'],3)) for k,v in d['roofline']['stages'].items()})
except Exception as e: print('no json', e); print(open("gpurun_out/k3_$wl.err").read()[-800:])
PY
  done
done
