#!/usr/bin/env python
"""profiles/rNN_workloads.md from the gpurun_out/w_*.json lines that tools/gpu_workloads.sh writes.  usage: make_workloads_md.py r01"""
import json
import os
import sys

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
peak = json.load(open(os.path.join(root, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(root, "MEASURED_PEAKS.json")) else 6548.2
names = ["raft_realtime_640x480_b64_f16", "fastacv_640x480_b64_f32", "hitnet_640x480_b64_f32", "fastacv_kitti_1242x375_b256_f32", "stream_1080p_b64_f32",
         "stream_1080p_resize_640x480_b64_f32", "uhd_2160p_b16_f32"]
rows = []
for wl in names:
    p = os.path.join(root, "gpurun_out", f"w_{wl}.json")
    if not os.path.exists(p):
        continue
    d = json.load(open(p))
    a, b = d["roofline"]["stages"]["rectify"], d["roofline"]["stages"]["reproject"]
    rows.append(f"| {wl} | {round(d['value'])} | {d['ms_per_step']:.3f} | {a['ms_per_launch']:.4f} ({round(a['GBps'])}, {a['GBps'] / peak:.2f}) | "
                f"{b['ms_per_launch']:.4f} ({round(b['GBps'])}, {b['GBps'] / peak:.2f}) | {d['roofline']['whole_step']['frac']:.3f} | "
                f"{round(d['e2e']['value'])} ({d['e2e']['pcie_gbs_per_gpu']:.1f}) | {d['clocks']['sm_mhz']:.0f} {','.join(d['clocks']['reasons']) or '-'} |")
print(f"| workload | pairs/s (device) | ms/step | stage A ms (GB/s, frac) | stage B ms (GB/s, frac) | whole step frac | e2e pairs/s (PCIe GB/s) | SM MHz, throttle reasons |")
print("|---|---|---|---|---|---|---|---|")
print("\n".join(rows))
