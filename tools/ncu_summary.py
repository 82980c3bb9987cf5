#!/usr/bin/env python
"""Markdown table from an ncu report: `ncu_summary.py X.ncu-rep "section title" >> profiles/rNN_ncu_summary.md`.
One row per distinct kernel (first launch captured), the metrics the roofline discussion in DESIGN.md uses."""
import csv
import subprocess
import sys

rep, title = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
COLS = [("gpu__time_duration.sum", "gpu__time_duration"), ("dram__bytes_read.sum", "dram__bytes_read"), ("dram__bytes_write.sum", "dram__bytes_write"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput"), ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__issue_active"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "sm__warps_active"),
        ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "pipe_lsu"), ("launch__registers_per_thread", "registers_per_thread"),
        ("smsp__inst_executed.sum", "smsp__inst_executed"), ("lts__t_sector_hit_rate.pct", "lts__t_sector_hit_rate"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem_bank_conflicts")]
idx = {n: hdr.index(n) for n, _ in COLS if n in hdr}
ki = hdr.index("Kernel Name")
print(f"\n## {title}\n")
print("| kernel | " + " | ".join(short for n, short in COLS if n in idx) + " |")
print("|---|" + "---|" * len(idx))
seen = set()
for r in rows[2:]:
    name = r[ki].replace("void unnamed>::", "").split("(")[0]
    if name in seen:
        continue
    seen.add(name)
    cells = []
    for n, _ in COLS:
        if n not in idx:
            continue
        v, u = r[idx[n]], units[idx[n]]
        try:
            v = f"{float(v):.4g}"
        except ValueError:
            pass
        cells.append(f"{v} {u}".strip())
    print(f"| `{name}` | " + " | ".join(cells) + " |")
