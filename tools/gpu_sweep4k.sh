# 3840x2160, rectify + preprocess + reproject, batch sweep on one GPU
mkdir -p gpurun_out
echo "| N pairs | pairs/s | ms/step | stage A ms (GB/s) | stage B ms (GB/s) | whole-step frac |" > gpurun_out/sweep4k.md
echo "|---|---|---|---|---|---|" >> gpurun_out/sweep4k.md
for n in 1 2 4 8 16 32 64 128; do
python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-e2e --workload uhd_2160p_b16_f32 --batch $n 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); s=d['roofline']['stages']
print('| $n | %d | %.3f | %.4f (%d) | %.4f (%d) | %.3f |' % (d['value'], d['ms_per_step'], s['rectify']['ms_per_launch'], s['rectify']['GBps'], s['reproject']['ms_per_launch'], s['reproject']['GBps'], d['roofline']['whole_step']['frac']))" | tee -a gpurun_out/sweep4k.md
done
