# round-2 evidence run: GPU tests, smoke, default bench (both arms), heat map / compact steps, all workloads, launch list + full ncu captures
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_default.err
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref rc=$?"
python bench.py --heatmap --no-cpu-baseline --no-extra > gpurun_out/bench_heatmap.json 2> gpurun_out/bench_heatmap.err; echo "heatmap bench rc=$?"
python bench.py --reproject compact --no-cpu-baseline --no-extra > gpurun_out/bench_compact.json 2> gpurun_out/bench_compact.err; echo "compact bench rc=$?"
bash tools/gpu_workloads.sh
CMD="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-ref-kernels --no-extra"
$CMD > gpurun_out/r02_plain.json 2> gpurun_out/r02_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r02_launches_final.csv $CMD > gpurun_out/r02_ncu1.log 2>&1
echo "launch list rc=$?"
bash tools/gpu_prof2.sh r02ws rectify_ws
bash tools/gpu_prof2.sh r02rep reproject_dense
CMD2="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-ref-kernels --workload stream_1080p_resize_640x480_b64_f32 --opt fused_resize=1"
$CMD2 > gpurun_out/r02f_plain.json 2> gpurun_out/r02f_plain.err && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 60 --csv --log-file gpurun_out/r02_launches_fused_resize.csv $CMD2 > gpurun_out/r02f_ncu1.log 2>&1
echo "fused launch list rc=$?"
bash tools/gpu_r2i.sh | tail -8
python tools/jpeg_capture_bench.py --n 64 > gpurun_out/jpeg64.json 2> gpurun_out/jpeg64.err; echo "jpeg rc=$?"; cut -c1-300 gpurun_out/jpeg64.json
bash tools/gpu_prof3.sh r02heat heatmap_kernel --no-extra --heatmap
