#!/usr/bin/env python
"""bench.py - stereo pairs/s of the geometry hot path (rectify + preprocess + reproject) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

One "step" = one pass of the hot path over one batch of synthetic stereo pairs:
    sa_rectify_preprocess (fused rectify + normalise, both eyes, whole batch, one launch)
    -> [network slot: the disparity buffer was filled once by the synthetic stub before timing]
    -> sa_reproject dense (whole batch, one launch).
`value`      device-timed, inputs already resident in HBM (CUDA events, max over ranks);
`e2e`        the same metric through the reference-facing host API (sa_run_batch: pinned host images in, point cloud +
             disparity + rectified images out), H2D/D2H inside the timed region;
`roofline`   the dominant kernel's algorithmic bytes / its event-timed duration vs the measured HBM peak;
`cpu_baseline` the reference's CPU path (cv2 = the OpenCV calls the reference makes, maps recomputed per frame as it
             does, + the oracle's C/OpenMP restatement of its two CUDA kernels) on the host cores, bounded sample.
Frames are independent: with N GPUs every rank processes its own batch (weak scaling), no collective on the data path;
torch.distributed is used only for the start/stop barrier and the max-over-ranks of the elapsed time.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")

METRIC = "stereo_pairs_per_sec_rectify_preproc_reproject"
UNIT = "pairs/s"

# BASELINE.json configs (index -> workload).  configs[1] is the one the metric is quoted on at N=1.
WORKLOADS = {
    "raft_realtime_640x480_b64_f16": dict(model="RAFTStereo", W=640, H=480, batch=64, dtype="f16", calib=("calib_a", 1.0)),
    "fastacv_640x480_b64_f32": dict(model="FastACVNet_plus", W=640, H=480, batch=64, dtype="f32", calib=("calib_a", 1.0)),
    "hitnet_640x480_b64_f32": dict(model="HitNet", W=640, H=480, batch=64, dtype="f32", calib=("calib_b", 1.0)),
    "fastacv_kitti_1242x375_b256_f32": dict(model="FastACVNet_plus", W=1242, H=375, batch=256, dtype="f32", calib=("calib_a", 1242 / 640)),
    "stream_1080p_b64_f32": dict(model="FastACVNet_plus", W=1920, H=1080, batch=64, dtype="f32", calib=("calib_a", 3.0)),
    "uhd_2160p_b16_f32": dict(model="FastACVNet_plus", W=3840, H=2160, batch=16, dtype="f32", calib=("calib_a", 6.0)),
}
# not BASELINE configs: diagnostic shapes used while tuning the 2-byte-aligned (W % 4 != 0) path
WORKLOADS.update({
    # BASELINE configs[3] (ii): 1080p capture, network at 640x480 (rectify at full size -> 8-bit INTER_LINEAR resize -> tensors)
    "stream_1080p_resize_640x480_b64_f32": dict(model="FastACVNet_plus", W=1920, H=1080, net=(640, 480), batch=64, dtype="f32", calib=("calib_a", 3.0)),
    "exp_1240x375_b256_f32": dict(model="FastACVNet_plus", W=1240, H=375, batch=256, dtype="f32", calib=("calib_a", 1240 / 640)),
    "exp_1242x375_b64_f32": dict(model="FastACVNet_plus", W=1242, H=375, batch=64, dtype="f32", calib=("calib_a", 1242 / 640)),
    "exp_1242x376_b256_f32": dict(model="FastACVNet_plus", W=1242, H=376, batch=256, dtype="f32", calib=("calib_a", 1242 / 640)),
})
DEFAULT_WORKLOAD = "raft_realtime_640x480_b64_f16"


def bytes_per_pair(wl, stage):
    """Algorithmic HBM bytes per stereo pair (SURVEY.md section 8d / DESIGN.md): maps, LUTs and Q are batch-shared
    constants and are not counted."""
    WH = wl["W"] * wl["H"]                                     # full-resolution image
    wh = wl["net"][0] * wl["net"][1] if wl.get("net") else WH  # network / disparity / point-cloud size (= WH unless resizing)
    t = 2 if wl["dtype"] == "f16" else 4
    # raw u8 in, rectified u8 out, tensors out (both eyes).  With a resize the intermediate traffic (rectified image read
    # back, resized u8 written and read) is NOT counted: it is not compulsory, a fully fused path would not have it.
    a = 2 * 3 * WH + 2 * 3 * WH + 2 * 3 * wh * t
    b = 4 * wh + 3 * wh + 24 * wh + (4 * wh if wl["model"] == "RAFTStereo" else 0)   # disparity, left u8, points (+ write-back)
    return {"rectify": a, "reproject": b, "total": a + b}[stage]


def make_inputs(wl, n, rank=0):
    """the demo frames plus per-frame uniform noise of +-8 (seed = frame index); other sizes:
    uniform random images (seed 2000 + i)."""
    from oracle import oracle as O
    W, H = wl["W"], wl["H"]
    L = np.empty((n, H, W, 3), np.uint8)
    R = np.empty((n, H, W, 3), np.uint8)
    if (W, H) == (640, 480):
        import cv2
        l0 = cv2.imread(os.path.join(GOLDEN, "left0.png"))
        r0 = cv2.imread(os.path.join(GOLDEN, "right0.png"))
        for i in range(n):
            rng = np.random.default_rng(rank * 100003 + i)
            L[i] = np.clip(l0.astype(np.int16) + rng.integers(-8, 9, l0.shape), 0, 255)
            R[i] = np.clip(r0.astype(np.int16) + rng.integers(-8, 9, r0.shape), 0, 255)
    else:
        base_l, base_r = O.synthetic_image(W, H, 2000 + rank), O.synthetic_image(W, H, 3000 + rank)
        for i in range(n):          # cheap per-frame variation: roll the random base image
            L[i] = np.roll(base_l, 17 * i, axis=1)
            R[i] = np.roll(base_r, 17 * i, axis=1)
    return L, R


def load_calib(wl):
    from oracle import oracle as O
    name, s = wl["calib"]
    c = O.load_calibration_yaml(os.path.join(GOLDEN, name + ".yml"))
    if s != 1.0:
        c = O.scale_calibration(c, s, crop_y=(480 * s - wl["H"]) / 2)
    return c


# ---------------------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons while the GPU is under load (B200_PROFILING.md clocks line)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        self.samples = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50", "-i", str(gpu_index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.proc.terminate()
        rows = [s for (t, s) in self.samples if t0 <= t <= t1] or [s for (_, s) in self.samples]
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            f = [x.strip() for x in r.split(",")]
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); power.append(float(f[2]))
            except (ValueError, IndexError):
                continue
            for nme, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


def bind_to_gpu_numa_node(gpu_index):
    """Pin this process to the CPUs of the NUMA node the GPU hangs off, so that the pinned host buffers of the end-to-end
    leg are allocated next to the GPU's PCIe root (matters with several ranks on a two-socket box).  Best effort."""
    try:
        bdf = subprocess.check_output(["nvidia-smi", "--query-gpu=pci.bus_id", "--format=csv,noheader", "-i", str(gpu_index)], text=True).strip().lower()
        bdf = bdf[-12:] if len(bdf) > 12 else bdf                      # 00000000:1B:00.0 -> 0000:1b:00.0
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return node
    except Exception:
        pass
    return None


def measured_hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---------------------------------------------------------------------------------------------------------------
def cpu_reference_pairs(wl, calib, L, R, D, threads, budget_s=None, max_pairs=None, cached_maps=False):
    """The reference's CPU path for this workload, pair by pair, as the reference runs it (batch 1):
    RectifyImage = 2x initUndistortRectifyMap + 2x remap EVERY frame (FastACVNet_plus_Algorithm.cpp:132-145) through cv2
    (the same OpenCV functions; falls back to the oracle's C restatement if cv2 is missing), then the C/OpenMP restatement
    of the reference's preprocess and reproject kernels (which have no CPU code in the reference).  Returns (pairs, seconds)."""
    from oracle import oracle as O
    from oracle import oracle_c as OC
    OC.set_threads(threads)
    try:
        import cv2
        cv2.setNumThreads(threads)
    except ImportError:
        cv2 = None
    mid = {"RAFTStereo": O.RAFTSTEREO, "HitNet": O.HITNET, "CREStereo": O.CRESTEREO, "FastACVNet_plus": O.FASTACVNET_PLUS}[wl["model"]]
    size = (wl["W"], wl["H"])
    keys = (("intrinsic_left", "distCoeffs_left", "R_L", "P1"), ("intrinsic_right", "distCoeffs_right", "R_R", "P2"))
    maps = [None, None]
    done, t0 = 0, time.perf_counter()
    while True:
        i = done % len(L)
        imgs = (L[i], R[i])
        rect = []
        for e in range(2):
            if maps[e] is None or not cached_maps:
                args = [calib[k] for k in keys[e]]
                maps[e] = cv2.initUndistortRectifyMap(*args, size, cv2.CV_16SC2) if cv2 else OC.init_maps(*args, size)
            rect.append(cv2.remap(imgs[e], maps[e][0], maps[e][1], cv2.INTER_LINEAR) if cv2 else OC.remap(imgs[e], *maps[e]))
        OC.preprocess_eye(mid, rect[0])
        OC.preprocess_eye(mid, rect[1])
        OC.reproject(mid, D[i % len(D)], rect[0], calib["Q"])
        done += 1
        el = time.perf_counter() - t0
        if (max_pairs and done >= max_pairs) or (budget_s and el >= budget_s):
            return done, el, ("cv2 " + cv2.__version__) if cv2 else "oracle_c"


def run_reference_arm(args, wl, wname):
    """--impl reference: rank 0 only, CPU, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle as O
    threads = os.cpu_count() or 1
    calib = load_calib(wl)
    sample = 8 if wl["W"] * wl["H"] <= 640 * 480 else 2
    L, R = make_inputs(wl, sample)
    D = [O.synthetic_disparity(wl["W"], wl["H"], 30 + i) * (-1 if wl["model"] == "RAFTStereo" else 1) for i in range(sample)]
    for _ in range(args.warmup):
        cpu_reference_pairs(wl, calib, L, R, D, threads, max_pairs=sample)
    t = 0.0
    for _ in range(args.steps):
        n, el, how = cpu_reference_pairs(wl, calib, L, R, D, threads, max_pairs=sample)
        t += el
    value = args.steps * sample / t
    desc = f"{sample} pairs per step of {wname} (batch-1 calls, maps recomputed every frame as the reference does; {how} for rectification, " \
           f"oracle C/OpenMP for preprocess+reproject, f32 tensors)"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8+f32",
            "data": "synthetic", "config": {"workload": wname, "pairs_per_step": sample, "size": [wl["W"], wl["H"]], "model": wl["model"]},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": desc},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=list(WORKLOADS))
    ap.add_argument("--batch", type=int, default=0, help="override the per-GPU batch of the workload")
    ap.add_argument("--e2e-steps", type=int, default=0, help="steps of the host-to-host leg (default min(steps, 20))")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--heatmap", action="store_true", help="include stage C in the step")
    ap.add_argument("--frame-fastest", type=int, default=-1)
    ap.add_argument("--opt", action="append", default=[], help="library tuning option key=value (sa_set_option), repeatable")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    wname = args.workload
    wl = dict(WORKLOADS[wname])
    if args.batch:
        wl["batch"] = args.batch

    if args.impl == "reference":
        run_reference_arm(args, wl, wname)
        return

    import torch
    import torch.distributed as dist
    from stereoalgorithms_b200 import _lib as SL
    from stereoalgorithms_b200.geometry import DENSE, F16, F32, PinnedArray, StereoGeometry

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    numa = bind_to_gpu_numa_node(local) if world > 1 else None
    if world > 1:
        # The data path has no collective (frames share nothing): the process group only carries the barrier around the timed
        # region and two scalar reductions (MAX of the elapsed time, SUM of the pairs), so gloo over loopback is enough and
        # keeps NCCL's start-up (and its version banner on stdout) out of a run that never moves data between GPUs.
        dist.init_process_group("gloo")
    n = wl["batch"]
    W, H = wl["W"], wl["H"]
    calib = load_calib(wl)
    L, R = make_inputs(wl, n, rank)
    dL, dR = torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()

    g = StereoGeometry(W, H, model=wl["model"], dtype=F16 if wl["dtype"] == "f16" else F32, max_batch=n, gpu_id=local, flags=SL.FLAG_TIMING,
                       net_size=wl.get("net"))
    w, h = wl.get("net") or (W, H)
    g.set_calibration(calib)
    if args.frame_fastest >= 0:
        g.set_option("frame_fastest", args.frame_fastest)
    for kv in args.opt:
        k_, v_ = kv.split("=")
        g.set_option(k_, int(v_))
    g.stub_disparity(n, seed=1234 + rank)
    g.synchronize()

    stream = torch.cuda.Stream()                              # a real (non-null) stream: NULL would mean "the handle's own stream"
    torch.cuda.set_stream(stream)
    sp = stream.cuda_stream                                   # every kernel of the step is launched on this stream
    assert sp != 0

    def step():
        g.rectify_preprocess(dL, dR, n=n, stream=sp)
        g.reproject(n, DENSE, to_host=False, stream=sp)
        if args.heatmap:
            g.heatmap(n, to_host=False, stream=sp)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local) if rank == 0 else None
    for _ in range(args.warmup):
        step()
    barrier()
    g.reset_timing()
    launches0 = g.launch_count
    t_wall0 = time.time()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    ev0.record(stream)
    for _ in range(args.steps):
        step()
    ev1.record(stream)
    barrier()
    elapsed = time.perf_counter() - t0                        # host clock, for information
    step_time = ev0.elapsed_time(ev1) / 1e3                   # device time of exactly K steps, events on the launch stream
    launches = g.launch_count - launches0
    rect_ms, rect_cnt = g.stage_time(SL.STAGE_RECTIFY)
    resize_ms = 0.0
    if wl.get("net"):
        resize_ms = g.stage_time(SL.STAGE_RESIZE)[0]         # resize + preprocess kernels
        rect_ms += resize_ms                                  # stage A = rectify + resize + preprocess
    rep_ms, rep_cnt = g.stage_time(SL.STAGE_REPROJECT)
    heat_ms, _ = g.stage_time(SL.STAGE_HEATMAP) if args.heatmap else (0.0, 0)
    dev_ms = rect_ms + rep_ms + heat_ms                      # sum of the per-kernel event pairs
    from stereoalgorithms_b200 import sharding
    pairs_total, step_time, launches = sharding.aggregate(n * args.steps, step_time, launches)     # SUM pairs, MAX time, SUM launches
    _, dev_s, _ = sharding.aggregate(0, dev_ms / 1e3)
    value = pairs_total / step_time

    # ---- e2e: host buffers through the reference-facing call --------------------------------------------------------
    e2e = None
    if not args.no_e2e:
        k2 = args.e2e_steps or min(args.steps, 20)
        npix = w * h
        hin = [PinnedArray(L.shape, np.uint8), PinnedArray(R.shape, np.uint8)]
        hin[0].array[:] = L
        hin[1].array[:] = R
        hout = [PinnedArray((n, npix, 6), np.float32), PinnedArray((n, h, w), np.float32), PinnedArray(L.shape, np.uint8), PinnedArray(R.shape, np.uint8)]

        def e2e_step():
            g.run_batch(hin[0].array, hin[1].array, n, rectify=True, stub_seed=0, pointcloud=hout[0].array, disparity=hout[1].array,
                        rect_left=hout[2].array, rect_right=hout[3].array)

        for _ in range(3):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(k2):
            e2e_step()                                        # synchronous: returns when the results are in host memory
        el2 = time.perf_counter() - t0
        pairs2, t2, _ = sharding.aggregate(n * k2, el2)
        h2d = 2 * n * W * H * 3
        d2h = n * (npix * (24 + 4) + W * H * 6)
        e2e = {"value": pairs2 / t2, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": k2,
               "api": "sa_run_batch (pinned host images in; point cloud + disparity + rectified L/R out)",
               "pcie_gbs_per_gpu": (h2d + d2h) * k2 / t2 / 1e9}
    t_wall2 = time.time()
    clocks = sampler.stop(t_wall0, t_wall2) if sampler else None

    if rank == 0:
        peak, peak_src = measured_hbm_peak()
        stages = {"rectify": (rect_ms, rect_cnt), "reproject": (rep_ms, rep_cnt)}
        dom = max(stages, key=lambda k: stages[k][0])
        dom_ms, dom_cnt = stages[dom]
        per_launch_bytes = bytes_per_pair(wl, dom) * n
        achieved = per_launch_bytes / (dom_ms / dom_cnt * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": "rectify_batch_kernel (fused rectify+preprocess)" if dom == "rectify" else "reproject_dense_kernel",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src, "traffic": None,
                "algorithmic_bytes_per_launch": per_launch_bytes, "avg_launch_ms": dom_ms / dom_cnt,
                "stages": {k: {"ms_per_launch": v[0] / max(v[1], 1), "GBps": bytes_per_pair(wl, k) * n / (v[0] / max(v[1], 1) * 1e-3) / 1e9,
                               "bytes_per_pair": bytes_per_pair(wl, k)} for k, v in stages.items()},
                "resize_preprocess_ms_per_step": resize_ms / max(rect_cnt, 1) if wl.get("net") else None,
                "whole_step": {"GBps": bytes_per_pair(wl, "total") * n * args.steps / step_time / 1e9,
                               "frac": bytes_per_pair(wl, "total") * n * args.steps / step_time / 1e9 / peak}}
        prof = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(prof):
            try:
                roof["traffic"] = json.load(open(prof)).get(wname, {}).get(dom)
            except Exception:
                pass
        cpu = None
        if not args.no_cpu_baseline and world == 1:
            from oracle import oracle as O
            threads = os.cpu_count() or 1
            m = min(n, 16)
            D = [O.synthetic_disparity(W, H, 30 + i) * (-1 if wl["model"] == "RAFTStereo" else 1) for i in range(4)]
            cpu_reference_pairs(wl, calib, L[:m], R[:m], D, threads, max_pairs=4)
            pairs, secs, how = cpu_reference_pairs(wl, calib, L[:m], R[:m], D, threads, budget_s=12.0)
            pairs_c, secs_c, _ = cpu_reference_pairs(wl, calib, L[:m], R[:m], D, threads, budget_s=4.0, cached_maps=True)
            cpu = {"value": pairs / secs, "unit": UNIT, "cores": threads, "kind": "port",
                   "sample": f"{pairs} pairs of {wname} in {secs:.1f} s, batch-1 calls, maps recomputed every frame as the reference does "
                             f"({how} for initUndistortRectifyMap+remap, oracle C/OpenMP for preprocess+reproject, f32 tensors)",
                   "maps_cached_value": pairs_c / secs_c}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": 1e3 * step_time / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u8 fixed-point + f32 (%s tensors)" % wl["dtype"], "data": "synthetic",
                "config": {"workload": wname, "model": wl["model"], "size": [W, H], "net_size": [w, h], "pairs_per_gpu_per_step": n, "tensor_dtype": wl["dtype"],
                           "reproject": "dense", "heatmap": bool(args.heatmap), "parallelism": f"frame-sharded x{world}, no collective",
                           "numa_node_rank0": numa,
                           "l2": "working set per step %.0f MB >> 126 MB L2 (no flush needed)" % (bytes_per_pair(wl, "total") * n / 1e6),
                           "network_slot": "synthetic disparity written once before timing (RAFTStereo: its sign alternates per step, same work)"},
                "kernel_ms_per_step": 1e3 * dev_s / args.steps, "host_ms_per_step": 1e3 * elapsed / args.steps,
                "e2e": e2e, "gpu_launches": launches, "roofline": roof, "cpu_baseline": cpu, "clocks": clocks}
        print(json.dumps(line))
    g.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
