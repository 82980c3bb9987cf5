mkdir -p gpurun_out
N=${1:-2}
nvidia-smi -L
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 100 --warmup 10 > gpurun_out/bench_g$N.json 2> gpurun_out/bench_g$N.err; echo "rc=$?"; cat gpurun_out/bench_g$N.json; tail -5 gpurun_out/bench_g$N.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/bench_ref_g$N.json 2> gpurun_out/bench_ref_g$N.err; echo "ref rc=$?"; cat gpurun_out/bench_ref_g$N.json | cut -c1-300
