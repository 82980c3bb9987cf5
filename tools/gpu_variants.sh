for flags in "$@"; do
  echo "=== variant: [$flags]"
  SA_EXTRA_NVCC_FLAGS="$flags" python stereoalgorithms_b200/build.py --force > /dev/null 2>&1 || { echo build failed; continue; }
  for wl in raft_realtime_640x480_b64_f16 fastacv_kitti_1242x375_b256_f32 stream_1080p_b64_f32; do
    python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-e2e --workload $wl 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$wl', round(d['value']), {k:(round(v['ms_per_launch'],4), round(v['GBps'])) for k,v in d['roofline']['stages'].items()})"
  done
done
python stereoalgorithms_b200/build.py --force > /dev/null 2>&1
