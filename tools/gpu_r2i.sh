mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-ref-kernels --no-extra --reproject compact"
$CMD > gpurun_out/cmp_plain.json 2> gpurun_out/cmp_plain.err && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum --clock-control none -c 60 --csv --log-file gpurun_out/r02_launches_compact.csv $CMD > gpurun_out/cmp_ncu.log 2>&1
echo "rc=$?"
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/r02_launches_compact.csv')) if len(r)>10]
hdr=rows[0]; ki=hdr.index('Kernel Name'); mi=hdr.index('Metric Name'); vi=hdr.index('Metric Value'); ii=hdr.index('ID')
d={}
for r in rows[1:]:
    d.setdefault((r[ii], r[ki].split('(')[0][-40:]),{})[r[mi]]=r[vi]
for k,v in list(d.items())[-12:]: print(k, v)
PY
