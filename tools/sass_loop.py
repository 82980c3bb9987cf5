#!/usr/bin/env python
"""Instruction mix of the frame loop of one rectify kernel in libstereo_geom.so (cuobjdump -sass): finds the backward branch that
contains a BAR.SYNC and the most dot-product instructions and prints the opcode histogram (+ the listing with -l).
    python tools/sass_loop.py 'rectify_batch16_kernelILi1ELb1ELb1ELi640ELi480E' [-l]"""
import collections
import os
import re
import subprocess
import sys

so = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "stereoalgorithms_b200", "libstereo_geom.so")
pat = sys.argv[1]
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
ins, on = [], False
for l in txt.splitlines():
    if "Function :" in l:
        if on:
            break
        on = pat in l
        continue
    if on:
        m = re.match(r"\s*/\*([0-9a-f]{4,5})\*/\s+(.*?);", l)
        if m:
            ins.append((int(m.group(1), 16), m.group(2)))
best, best_key = None, None
for a, t in ins:
    m = re.search(r"BRA(?:\.U)?\s+(?:!?U?P\d+,\s*)?(0x[0-9a-f]+)", t)
    if m:
        tgt = int(m.group(1), 16)
        if tgt < a:
            body = [(x, y) for x, y in ins if tgt <= x <= a]
            nidp = sum("IDP" in y for _, y in body)              # the blend loop is the one full of dot products
            if any("BAR.SYNC" in y for _, y in body) and (best is None or (nidp, -len(body)) > best_key):
                best, best_key = body, (nidp, -len(body))
print("total instructions:", len(ins))
if best:
    c = collections.Counter(re.sub(r"^@!?U?P\d+\s+", "", t).split()[0].split(".")[0] for _, t in best)
    nbar = sum("BAR.SYNC" in t for _, t in best)
    print(f"loop {hex(best[0][0])}..{hex(best[-1][0])}: {len(best)} instructions, {nbar} barrier(s) (= frames per iteration)")
    print(sorted(c.items(), key=lambda kv: -kv[1]))
    if "-l" in sys.argv:
        for a, t in best:
            print(hex(a), t)
