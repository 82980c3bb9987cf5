# one 8-GPU call: host copy ceiling, BASELINE configs[2]-[4] across GPUs, bench.py at N = 8 / 4 (torchrun, as the driver launches it)
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1; nproc >> gpurun_out/topo.txt; free -g >> gpurun_out/topo.txt; numactl -H >> gpurun_out/topo.txt 2>&1
( time timeout 300 python tools/pcie_ceiling.py --gpus 1,2,4,8 --seconds 0.6 --json gpurun_out/pcie_ceiling.json --md gpurun_out/r02_pcie_ceiling.md > gpurun_out/pcie.log 2>&1 ) 2>&1 | grep real; echo "pcie rc=$?"
( time timeout 600 python tools/scaling_configs.py --gpus 1,2,4,8 --md gpurun_out/r02_scaling_configs.md > gpurun_out/scaling.log 2>&1 ) 2>&1 | grep real; echo "scaling rc=$?"; tail -3 gpurun_out/scaling.log
for n in 8 4 4; do
( time timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n --steps 20 --warmup 5 --no-extra > gpurun_out/bench_${n}gpu_$RANDOM.json 2> gpurun_out/bench_${n}gpu.err ) 2>&1 | grep real; echo "bench $n rc=$?"
done
