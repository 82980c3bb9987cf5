mkdir -p gpurun_out
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 7 --print-limit 20 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -q -x -k "rectified or tensors or reproject or heatmap or resize or noalloc or side_by_side" > gpurun_out/memcheck.log 2>&1
echo "memcheck rc=$?"
grep -E "ERROR SUMMARY|Invalid|passed|failed|error" gpurun_out/memcheck.log | head -20
