#!/usr/bin/env python
"""Key counters of the first kernel in each ncu report, side by side: python tools/ncu_stalls.py a.ncu-rep b.ncu-rep ..."""
import csv
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum"]
cols = {}
for rep in sys.argv[1:]:
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, r = rows[0], rows[2]
    d = dict(zip(hdr, r))
    d["_name"] = d.get("Kernel Name", "")[:70]
    cols[rep] = d
names = list(cols)
print("kernel", [cols[n]["_name"] for n in names])
keys = WANT + sorted(k for k in cols[names[0]] if "warp_issue_stalled" in k and k.endswith("_per_warp_active.pct"))
for k in keys:
    vals = [cols[n].get(k, "-") for n in names]
    def f(v):
        try:
            return f"{float(v.replace(',', '')):.5g}"
        except ValueError:
            return v
    print(f"{k:95s}", *[f"{f(v):>14s}" for v in vals])
