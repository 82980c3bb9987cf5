mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --workload fastacv_kitti_1242x375_b256_f32 --batch 64"
$CMD > gpurun_out/k_plain.json 2> gpurun_out/k_plain.err && \
ncu --set full --clock-control none --import-source on -k regex:'rectify_batch_kernel' -s 2 -c 1 -o gpurun_out/k_full $CMD > gpurun_out/k_ncu.log 2>&1
echo rc=$?
