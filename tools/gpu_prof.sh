# usage: bash tools/gpu_prof.sh <tag> [bench args...]   -> gpurun_out/<tag>_*.{json,csv,ncu-rep}
tag=$1; shift
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline $*"
$CMD > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu1.log 2>&1
echo "launch list rc=$?"
$CMD > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"${SA_PROF_KERNEL:-rectify_batch_kernel}" -s 2 -c 2 -o gpurun_out/${tag}_full $CMD > gpurun_out/${tag}_ncu2.log 2>&1
echo "full capture rc=$?"
ls -la gpurun_out | tail -12
