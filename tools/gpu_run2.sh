mkdir -p gpurun_out
python bench.py --steps 200 --warmup 20 > gpurun_out/bench2.json 2> gpurun_out/bench2.err; echo "bench rc=$?"; cat gpurun_out/bench2.json | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print({k:d[k] for k in ('value','ms_per_step','kernel_ms_per_step','host_ms_per_step')}); print(d['roofline']['stages']); print(d['e2e'])"
bash tools/gpu_prof.sh r1
