# round 2, first evidence run: GPU tests, default bench (headline + extras), reference arm, copy ceiling on one GPU
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
( time python bench.py --steps 20 --warmup 5 > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err ) 2>&1 | grep real; echo "bench rc=$?"; tail -3 gpurun_out/bench_default.err
( time python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err ) 2>&1 | grep real; tail -3 gpurun_out/bench_reference.err
python bench.py --steps 50 --warmup 5 --no-extra --no-cpu-baseline --no-e2e --heatmap > gpurun_out/bench_heatmap.json 2> gpurun_out/bench_heatmap.err; echo "heatmap rc=$?"
python bench.py --steps 50 --warmup 5 --no-extra --no-cpu-baseline --no-e2e --reproject compact > gpurun_out/bench_compact.json 2> gpurun_out/bench_compact.err; echo "compact rc=$?"
python tools/pcie_ceiling.py --gpus 1 --seconds 0.7 --json gpurun_out/pcie_1gpu.json > gpurun_out/pcie_1gpu.log 2>&1; echo "pcie rc=$?"
