#!/usr/bin/env python
"""What can this box move between pinned host memory and its GPUs?  (the end-to-end figure stops
scaling at ~100 GB/s through the host; is that the library or the box?)

Bare cudaMemcpyAsync loops in 64 MB blocks from/to cudaMallocHost memory, nothing else on the GPU:
    h2d      host -> device only
    d2h      device -> host only
    bidir    both directions at once on two streams (what sa_run_batch's side streams do)
    mix      the traffic mix of the headline step: 15 % H2D + 85 % D2H (118 MB in, 668 MB out per 64 pairs of 640x480)
for G = 1, 2, 4, 8 GPUs at once, in two process models:
    procs    one process per GPU (what torchrun / bench.py --gpus N does)
    threads  one process, one host thread per GPU (what sa_run_batch_multi does)

    python tools/pcie_ceiling.py [--gpus 1,2,4,8] [--seconds 1.0] [--md profiles/r02_pcie_ceiling.md]
"""
import argparse
import json
import os
import sys
import threading
import time

BLK = 64 << 20
NBUF = 8            # 512 MB per direction per GPU


def worker(dev, start, stop_after, out, idx):
    import torch
    torch.cuda.set_device(dev)
    hin = torch.empty(NBUF * BLK, dtype=torch.uint8).pin_memory()
    hout = torch.empty(NBUF * BLK, dtype=torch.uint8).pin_memory()
    hin.fill_(1)
    hout.fill_(0)                                             # touch every page before timing
    din = torch.empty(NBUF * BLK, dtype=torch.uint8, device="cuda")
    dout = torch.ones(NBUF * BLK, dtype=torch.uint8, device="cuda")
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    res = {}
    for mode in ("h2d", "d2h", "bidir", "mix"):
        torch.cuda.synchronize()
        start.wait()
        t0 = time.perf_counter()
        moved_in = moved_out = 0
        k = 0
        while time.perf_counter() - t0 < stop_after:
            b = (k % NBUF) * BLK
            if mode in ("h2d", "bidir") or (mode == "mix" and k % 7 == 0):
                with torch.cuda.stream(s_in):
                    din[b:b + BLK].copy_(hin[b:b + BLK], non_blocking=True)
                moved_in += BLK
            if mode in ("d2h", "bidir", "mix"):
                with torch.cuda.stream(s_out):
                    hout[b:b + BLK].copy_(dout[b:b + BLK], non_blocking=True)
                moved_out += BLK
            k += 1
            if k % 4 == 0:                                    # keep at most a few blocks queued per stream
                s_in.synchronize()
                s_out.synchronize()
        torch.cuda.synchronize()
        el = time.perf_counter() - t0
        res[mode] = {"h2d_gbs": moved_in / el / 1e9, "d2h_gbs": moved_out / el / 1e9}
    out[idx] = res


def proc_main(dev, barrier, seconds, q, idx):
    class B:                                                  # multiprocessing.Barrier with the threading.Barrier call shape
        def wait(self):
            barrier.wait()
    out = {}
    worker(dev, B(), seconds, out, idx)
    q.put((idx, out[idx]))


def run(model, G, seconds):
    if model == "threads":
        start = threading.Barrier(G)
        out = {}
        ts = [threading.Thread(target=worker, args=(g, start, seconds, out, g)) for g in range(G)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        return [out[g] for g in range(G)]
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    barrier, q = ctx.Barrier(G), ctx.Queue()
    ps = [ctx.Process(target=proc_main, args=(g, barrier, seconds, q, g)) for g in range(G)]
    for p in ps:
        p.start()
    got = dict(q.get(timeout=600) for _ in ps)
    for p in ps:
        p.join()
    return [got[g] for g in range(G)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", default="1,2,4,8")
    ap.add_argument("--seconds", type=float, default=1.0)
    ap.add_argument("--models", default="procs,threads")
    ap.add_argument("--md", default=None)
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    import torch
    have = torch.cuda.device_count()
    rows = []
    for model in args.models.split(","):
        for G in [int(x) for x in args.gpus.split(",")]:
            if G > have:
                continue
            per = run(model, G, args.seconds)
            for mode in ("h2d", "d2h", "bidir", "mix"):
                tot_in = sum(p[mode]["h2d_gbs"] for p in per)
                tot_out = sum(p[mode]["d2h_gbs"] for p in per)
                rows.append({"model": model, "gpus": G, "mode": mode, "h2d_total_gbs": tot_in, "d2h_total_gbs": tot_out,
                             "total_gbs": tot_in + tot_out, "per_gpu_gbs": (tot_in + tot_out) / G,
                             "min_gpu_gbs": min(p[mode]["h2d_gbs"] + p[mode]["d2h_gbs"] for p in per)})
                print(json.dumps(rows[-1]), flush=True)
    if args.json:
        json.dump(rows, open(args.json, "w"), indent=1)
    if args.md:
        with open(args.md, "w") as f:
            f.write("# Host <-> GPU copy ceiling of the B200 box (bare pinned cudaMemcpyAsync, 64 MB blocks)\n\n")
            f.write(f"`python tools/pcie_ceiling.py --gpus {args.gpus} --seconds {args.seconds}`; {have} GPU(s) visible, {os.cpu_count()} host CPUs.  GB/s = 1e9 bytes/s.\n\n")
            f.write("| process model | GPUs | mode | H2D total | D2H total | total | per GPU | slowest GPU |\n|---|---|---|---|---|---|---|---|\n")
            for r in rows:
                f.write(f"| {r['model']} | {r['gpus']} | {r['mode']} | {r['h2d_total_gbs']:.1f} | {r['d2h_total_gbs']:.1f} | {r['total_gbs']:.1f} | "
                        f"{r['per_gpu_gbs']:.1f} | {r['min_gpu_gbs']:.1f} |\n")


if __name__ == "__main__":
    sys.exit(main())
