# usage: bash tools/gpu_prof3.sh <tag> <kernel regex> [bench args...]  -> like gpu_prof2.sh without --no-extra being forced (pass what you need)
tag=$1; kre=$2; shift; shift
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-ref-kernels $*"
$CMD > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err && \
ncu --set full --clock-control none --import-source on -k regex:"$kre" -s 2 -c 1 -f -o gpurun_out/${tag}_full $CMD > gpurun_out/${tag}_ncu.log 2>&1
echo "$tag full capture rc=$?"
