# round-end evidence run: GPU tests, smoke, default bench (both arms), all workloads, launch list + full ncu capture
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_default.err
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref rc=$?"
python bench.py --heatmap --no-cpu-baseline > gpurun_out/bench_heatmap.json 2> gpurun_out/bench_heatmap.err; echo "heatmap bench rc=$?"
bash tools/gpu_workloads.sh
bash tools/gpu_prof.sh r8
SA_PROF_KERNEL=reproject_dense_kernel bash tools/gpu_prof.sh r8b
