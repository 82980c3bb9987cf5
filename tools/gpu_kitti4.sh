# full ncu captures of the batch kernel at 1242 (misaligned tensor rows) and 1248 (aligned), 64 pairs each, for the side-by-side in profiles/r02_ncu_summary.md
bash tools/gpu_prof3.sh k1242 rectify_batch --workload fastacv_kitti_1242x375_b256_f32 --batch 64 --opt bulk=0
bash tools/gpu_prof3.sh k1248 rectify_batch --workload exp_1248x375_b256_f32 --batch 64 --opt bulk=0
