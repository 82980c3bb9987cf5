#!/usr/bin/env python
"""Summarise `ncu -i X.ncu-rep --page source --csv`: per kernel, the SASS with executed-instruction counts and
stall samples.  usage: ncu_sass.py src.csv [kernel-substring] [--top N | --all]"""
import csv
import sys

path = sys.argv[1]
pat = sys.argv[2] if len(sys.argv) > 2 and not sys.argv[2].startswith("--") else ""
show_all = "--all" in sys.argv
top = 40
if "--top" in sys.argv:
    top = int(sys.argv[sys.argv.index("--top") + 1])
rows = list(csv.reader(open(path)))
kernels, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "ins": []}
        kernels.append(cur)
    elif r and r[0] == "Address":
        cur["hdr"] = r
    elif cur is not None and cur["hdr"] and len(r) == len(cur["hdr"]):
        cur["ins"].append(r)
seen = set()
for k in kernels:
    if pat not in k["name"] or k["name"] in seen:
        continue
    seen.add(k["name"])
    h = {n: i for i, n in enumerate(k["hdr"])}
    ins = k["ins"]
    tot_exec = sum(int(r[h["Instructions Executed"]]) for r in ins)
    tot_samp = sum(int(r[h["# Samples"]]) for r in ins)
    print("=" * 100)
    print(k["name"])
    print(f"SASS instructions: {len(ins)}   warp-instructions executed: {tot_exec}   samples: {tot_samp}")
    stalls = [n for n in k["hdr"] if n.startswith("stall_") and "Not Issued" not in n]
    agg = {s: sum(int(r[h[s]]) for r in ins) for s in stalls}
    print("stall totals:", ", ".join(f"{s[6:]}={v}" for s, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v))
    # opcode histogram weighted by executions
    ops = {}
    for r in ins:
        op = r[h["Source"]].split()[0] if not r[h["Source"]].strip().startswith("@") else r[h["Source"]].split()[1]
        op = op.split(".")[0]
        ops[op] = ops.get(op, 0) + int(r[h["Instructions Executed"]])
    print("opcode mix (% of executed):", ", ".join(f"{o}={100*v/tot_exec:.1f}" for o, v in sorted(ops.items(), key=lambda kv: -kv[1])[:25]))
    order = range(len(ins)) if show_all else sorted(range(len(ins)), key=lambda i: -int(ins[i][h["# Samples"]]))[:top]
    for i in (order if show_all else sorted(order)):
        r = ins[i]
        st = {s[6:]: int(r[h[s]]) for s in stalls if int(r[h[s]])}
        main = ",".join(f"{a}:{b}" for a, b in sorted(st.items(), key=lambda kv: -kv[1])[:3])
        print(f"{i:4d} ex={int(r[h['Instructions Executed']]):>9d} smp={int(r[h['# Samples']]):>6d}  {r[h['Source']].strip():<70s} {main}")
