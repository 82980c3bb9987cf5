# frames per CTA (option batch_ctas = CTAs per SM the chunking aims at; 0 = automatic ~128) on the stage A workloads: does a longer CTA (smaller start-up share) pay?
mkdir -p gpurun_out
for wl in stream_1080p_b64_f32 raft_realtime_640x480_b64_f16; do
for v in 0 4 8 16 32 64; do
    python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-e2e --no-ref-kernels --workload $wl --opt batch_ctas=$v > gpurun_out/h_$v.json 2> gpurun_out/h_$v.err
    python - <<PY
import json
try:
    d=json.load(open("gpurun_out/h_$v.json")); print("$wl batch_ctas=$v", {k:(round(x['ms_per_launch'],4), round(x['frac'],3)) for k,x in d['roofline']['stages'].items() if k=='rectify'})
except Exception as e: print('no json', e); print(open("gpurun_out/h_$v.err").read()[-800:])
PY
done
done
