# run-to-run spread of the default bench line (headline + extra workloads): five fresh processes
mkdir -p gpurun_out
for i in 1 2 3 4 5; do python bench.py --no-cpu-baseline --no-ref-kernels > gpurun_out/rep_$i.json 2> gpurun_out/rep_$i.err; done
python - <<'PY'
import json
rows=[]
for i in range(1,6):
    d=json.load(open(f"gpurun_out/rep_{i}.json"))
    r=d['roofline']['stages']
    w=d.get('workloads',{})
    rows.append((d['value'], d['ms_per_step'], r['rectify']['ms_per_launch'], r['rectify']['frac'], r['reproject']['ms_per_launch'], r['reproject']['frac'], d['e2e']['value'], d['e2e'].get('frac_of_host_ceiling'),
                 w.get('stream_1080p_b64_f32',{}).get('value'), w.get('fastacv_640x480_b64_f32',{}).get('value'), w.get('stream_1080p_resize_640x480_b64_f32',{}).get('value'), d['clocks']['sm_mhz'], ','.join(d['clocks']['reasons']) or '-'))
print("| run | pairs/s | ms/step | stage A ms (frac) | stage B ms (frac) | e2e pairs/s (of host ceiling) | 1080p pairs/s | 640x480 f32 pairs/s | 1080p->640x480 pairs/s | SM MHz, reasons |")
print("|---|---|---|---|---|---|---|---|---|---|")
for i,r in enumerate(rows,1):
    print(f"| {i} | {r[0]:,.0f} | {r[1]:.4f} | {r[2]:.4f} ({r[3]:.3f}) | {r[4]:.4f} ({r[5]:.3f}) | {r[6]:,.0f} ({r[7]:.3f}) | {r[8]:,.0f} | {r[9]:,.0f} | {r[10]:,.0f} | {r[11]:.0f} {r[12]} |")
PY
