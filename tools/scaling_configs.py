#!/usr/bin/env python
"""BASELINE configs[2]-[4] across the GPUs of one box, from ONE process (one host thread per GPU, no collective):

    kitti    FastACVNet_plus 1242x375, batch 256 sharded frame-wise over G GPUs           (configs[2])
    stream   1920x1080, batch 512, rectify + preprocess + reproject + heat map,           (configs[3])
             host-to-host with every output downloaded (sa_run_batch_multi)
    sweep4k  3840x2160, rectify + reproject, batch 1..1024 over G GPUs                    (configs[4])

Strong scaling: the batch is fixed, shard g = sa_shard_range(batch, g, G).  Device-timed = every GPU runs stage A + stage B
(+ C) on its resident shard, CUDA events per GPU, the slowest GPU counts.  e2e = sa_run_batch_multi on pinned host buffers.

    python tools/scaling_configs.py [--gpus 1,2,4,8] [--what kitti,stream,sweep4k] [--md profiles/r02_scaling_configs.md]
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PEAK = 6548.2


def calib_for(W, H):
    from oracle import oracle as O
    c = O.load_calibration_yaml(os.path.join(ROOT, "tests", "golden", "calib_a.yml"))
    s = W / 640
    return O.scale_calibration(c, s, crop_y=(480 * s - H) / 2) if s != 1.0 else c


def shard(n, g, G):
    per = -(-n // G)
    b = min(n, g * per)
    return b, min(n, b + per)


def device_timed(W, H, model, dtype, batch, G, heatmap=False, steps=10, warmup=3, flags=0):
    """-> (pairs/s, ms per step of the slowest GPU, per-stage bytes moved per pair)"""
    import torch
    from stereoalgorithms_b200 import _lib as SL
    from stereoalgorithms_b200.geometry import DENSE, F16, F32, StereoGeometry
    calib = calib_for(W, H)
    res = [None] * G
    start = threading.Barrier(G)
    err = []

    def worker(g):
        try:
            b, e = shard(batch, g, G)
            n = e - b
            if n == 0:
                start.wait(); start.wait()
                res[g] = 0.0
                return
            torch.cuda.set_device(g)
            geo = StereoGeometry(W, H, model=model, dtype=F16 if dtype == "f16" else F32, max_batch=n, gpu_id=g, flags=flags)
            geo.set_calibration(calib)
            gen = torch.Generator(device=f"cuda:{g}").manual_seed(1000 + g)
            L = torch.randint(0, 256, (n, H, W, 3), dtype=torch.uint8, device=f"cuda:{g}", generator=gen)
            R = torch.randint(0, 256, (n, H, W, 3), dtype=torch.uint8, device=f"cuda:{g}", generator=gen)
            geo.stub_disparity(n, seed=7 + g)
            geo.synchronize()
            st = torch.cuda.Stream(device=g)
            sp = st.cuda_stream

            def step():
                geo.rectify_preprocess(L, R, n=n, stream=sp)
                geo.reproject(n, DENSE, to_host=False, stream=sp)
                if heatmap:
                    geo.heatmap(n, to_host=False, stream=sp)
            for _ in range(warmup):
                step()
            torch.cuda.synchronize(g)
            start.wait()
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record(st)
            for _ in range(steps):
                step()
            ev1.record(st)
            torch.cuda.synchronize(g)
            start.wait()
            res[g] = ev0.elapsed_time(ev1) / steps
            geo.close()
            del L, R
            torch.cuda.empty_cache()
        except Exception as ex:                               # noqa: BLE001
            err.append(repr(ex))
            try:
                start.abort()
            except Exception:
                pass

    ts = [threading.Thread(target=worker, args=(g,)) for g in range(G)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    if err:
        return None, err[0][:160]
    ms = max(res)
    return batch / (ms * 1e-3), ms


def e2e_multi(W, H, model, dtype, batch, G, heatmap=False, steps=3):
    from stereoalgorithms_b200.geometry import F16, F32, PinnedArray, StereoGeometry, run_batch_multi
    calib = calib_for(W, H)
    per = -(-batch // G)
    geos = [StereoGeometry(W, H, model=model, dtype=F16 if dtype == "f16" else F32, max_batch=per, gpu_id=g) for g in range(G)]
    for g in geos:
        g.set_calibration(calib)
    npix = W * H
    L, R = PinnedArray((batch, H, W, 3), np.uint8), PinnedArray((batch, H, W, 3), np.uint8)
    rng = np.random.default_rng(5)
    base = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    for i in range(batch):
        L.array[i] = base
        R.array[i] = base
    out = dict(pointcloud=PinnedArray((batch, npix, 6), np.float32), disparity=PinnedArray((batch, H, W), np.float32),
               rect_left=PinnedArray((batch, H, W, 3), np.uint8), rect_right=PinnedArray((batch, H, W, 3), np.uint8))
    if heatmap:
        out["heatmap"] = PinnedArray((batch, H, W, 3), np.uint8)
    kw = {k: v.array for k, v in out.items()}
    run_batch_multi(geos, L.array, R.array, batch, rectify=True, stub_seed=3, **kw)
    t0 = time.perf_counter()
    for _ in range(steps):
        run_batch_multi(geos, L.array, R.array, batch, rectify=True, stub_seed=3, **kw)
    el = (time.perf_counter() - t0) / steps
    nbytes = 2 * batch * npix * 3 + sum(v.array.nbytes for v in out.values())
    for g in geos:
        g.close()
    for a in [L, R] + list(out.values()):
        a.free()
    return batch / el, el * 1e3, nbytes / el / 1e9


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", default="1,2,4,8")
    ap.add_argument("--what", default="kitti,stream,sweep4k")
    ap.add_argument("--md", default=None)
    ap.add_argument("--stream-batch", type=int, default=512)
    ap.add_argument("--sweep", default="1,4,16,64,256,1024")
    args = ap.parse_args()
    import torch
    have = torch.cuda.device_count()
    Gs = [g for g in (int(x) for x in args.gpus.split(",")) if g <= have]
    rows = []

    def emit(r):
        rows.append(r)
        print(json.dumps(r), flush=True)

    if "kitti" in args.what:
        W, H, n = 1242, 375, 256
        bpp = 67 * W * H
        for G in Gs:
            v, ms = device_timed(W, H, "FastACVNet_plus", "f32", n, G)
            r = {"config": "configs[2] 1242x375 b256 FastACV f32", "gpus": G, "device_pairs_s": v, "ms_per_step": ms, "roofline_frac": v * bpp / 1e9 / (PEAK * G) if v else None}
            try:
                ev, ems, gbs = e2e_multi(W, H, "FastACVNet_plus", "f32", n, G)
                r.update({"e2e_pairs_s": ev, "e2e_ms": ems, "e2e_host_gbs": gbs})
            except Exception as ex:                           # noqa: BLE001
                r["e2e_error"] = repr(ex)[:160]
            emit(r)
    if "stream" in args.what:
        W, H, n = 1920, 1080, args.stream_batch
        bpp = 74 * W * H
        for G in Gs:
            if n / G > 512:
                continue
            v, ms = device_timed(W, H, "FastACVNet_plus", "f32", n, G, heatmap=True, steps=5)
            r = {"config": f"configs[3] 1920x1080 b{n} FastACV f32 + heat map", "gpus": G, "device_pairs_s": v, "ms_per_step": ms,
                 "roofline_frac": v * bpp / 1e9 / (PEAK * G) if v else None}
            if G == max(Gs):
                for nb in (n, n // 2, n // 4):
                    try:
                        ev, ems, gbs = e2e_multi(W, H, "FastACVNet_plus", "f32", nb, G, heatmap=True, steps=2)
                        r.update({"e2e_batch": nb, "e2e_pairs_s": ev, "e2e_ms": ems, "e2e_host_gbs": gbs})
                        break
                    except Exception as ex:                   # noqa: BLE001
                        r["e2e_error"] = repr(ex)[:160]
            emit(r)
    if "sweep4k" in args.what:
        from stereoalgorithms_b200 import _lib as SL
        W, H = 3840, 2160
        bpp = 67 * W * H
        for n in [int(x) for x in args.sweep.split(",")]:
            for G in Gs:
                if G > n or n / G > 200:                      # a shard of 200 4K pairs is ~115 GB of device buffers
                    continue
                v, ms = device_timed(W, H, "FastACVNet_plus", "f32", n, G, steps=3 if n >= 64 else 10, warmup=2, flags=SL.FLAG_NO_HEATMAP_BUF)
                emit({"config": "configs[4] 3840x2160 rectify+reproject", "batch": n, "gpus": G, "device_pairs_s": v, "ms_per_step": ms,
                      "roofline_frac": v * bpp / 1e9 / (PEAK * G) if isinstance(v, float) else None})
    if args.md:
        with open(args.md, "w") as f:
            f.write("# BASELINE configs[2]-[4] across GPUs (tools/scaling_configs.py, one process, one host thread per GPU, no collective)\n\n")
            f.write(f"{have} GPU(s) visible.  Strong scaling: fixed batch, frame-wise shards.  roofline = algorithmic bytes (67 WH per pair, 74 WH with the heat map) "
                    f"/ (G x {PEAK} GB/s measured copy peak).\n\n")
            f.write("| config | batch | GPUs | device-timed pairs/s | ms/step | of roofline | e2e pairs/s (host to host) | e2e GB/s through the host |\n|---|---|---|---|---|---|---|---|\n")
            for r in rows:
                def fmt(x, d=0):
                    return "-" if x is None or isinstance(x, str) else f"{x:,.{d}f}"
                f.write(f"| {r['config']} | {r.get('batch', r.get('e2e_batch', ''))} | {r['gpus']} | {fmt(r.get('device_pairs_s'))} | {fmt(r.get('ms_per_step'), 3)} | "
                        f"{fmt(r.get('roofline_frac'), 3)} | {fmt(r.get('e2e_pairs_s'))} | {fmt(r.get('e2e_host_gbs'), 1)} |\n")


if __name__ == "__main__":
    main()
