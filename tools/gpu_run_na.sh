mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/na_alone.log 2>&1; echo "alone rc=$?"; tail -3 gpurun_out/na_alone.log
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/na_all.log 2>&1; echo "all rc=$?"; tail -3 gpurun_out/na_all.log
