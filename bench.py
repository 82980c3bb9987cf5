#!/usr/bin/env python
"""bench.py - stereo pairs/s of the geometry hot path (rectify + preprocess + reproject) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

The headline workload is BASELINE configs[1]; a default run then also times EXTRA_WORKLOADS (1080p, the f32 shape, the
fused resize path) and puts them under "workloads" in the same JSON line, each with its own roofline and e2e.

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
    "exp_1248x375_b256_f32": dict(model="FastACVNet_plus", W=1248, H=375, batch=256, dtype="f32", calib=("calib_a", 1248 / 640)),
    "exp_1248x384_b256_f32": dict(model="FastACVNet_plus", W=1248, H=384, batch=256, dtype="f32", calib=("calib_a", 1248 / 640)),
    "exp_1240x375_b256_f32": dict(model="FastACVNet_plus", W=1240, H=375, batch=256, dtype="f32", calib=("calib_a", 1240 / 640)),
    "exp_1244x375_b256_f32": dict(model="FastACVNet_plus", W=1244, H=375, batch=256, dtype="f32", calib=("calib_a", 1244 / 640)),
    "exp_1242x375_b256_f16": dict(model="FastACVNet_plus", W=1242, H=375, batch=256, dtype="f16", calib=("calib_a", 1242 / 640)),
    "exp_1242x375_b64_f32": dict(model="FastACVNet_plus", W=1242, H=375, batch=64, dtype="f32", calib=("calib_a", 1242 / 640)),
    "exp_1242x376_b256_f32": dict(model="FastACVNet_plus", W=1242, H=376, batch=256, dtype="f32", calib=("calib_a", 1242 / 640)),
})
DEFAULT_WORKLOAD = "raft_realtime_640x480_b64_f16"
# The metric is quoted "@640x480 & 1080p": a default run (no --workload) times these after the headline and reports them under
# "workloads" in the same JSON line - the 1080p half of the metric, the f32 / LUT shape the reference itself runs, and the
# fused rectify+resize path (BASELINE configs[3] (ii)).
EXTRA_WORKLOADS = ["stream_1080p_b64_f32", "fastacv_640x480_b64_f32", "stream_1080p_resize_640x480_b64_f32"]


def bytes_per_pair(wl, stage):
    """Algorithmic HBM bytes per stereo pair (SURVEY.md section 8d / DESIGN.md): maps, LUTs and Q are batch-shared
    constants and are not counted."""
    WH = wl["W"] * wl["H"]                                     # full-resolution image
    wh = wl["net"][0] * wl["net"][1] if wl.get("net") else WH  # network / disparity / point-cloud size (= WH unless resizing)
    t = 2 if wl["dtype"] == "f16" else 4
    # raw u8 in, rectified u8 out, tensors out (both eyes).  With a resize the intermediate traffic (rectified image read
    # back, resized u8 written and read) is NOT counted: it is not compulsory, a fully fused path would not have it.
    a = 2 * 3 * WH + 2 * 3 * WH + 2 * 3 * wh * t
    if wl.get("net"):
        a += 3 * wh                                            # the resized LEFT image is written once: stage B's colours
    b = 4 * wh + 3 * wh + 24 * wh + (4 * wh if wl["model"] == "RAFTStereo" else 0)   # disparity, left u8, points (+ write-back)
    c = 4 * wh + 3 * wh                                        # stage C: disparity in, BGR heat map out
    return {"rectify": a, "reproject": b, "heatmap": c, "total": a + b, "total_heatmap": a + b + c}[stage]


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

    def stop(self):
        if self.proc:
            time.sleep(0.12)
            self.proc.terminate()

    def window(self, t0, t1):
        """Clocks and throttle reasons of the samples taken in [t0, t1] (one workload: its warm-up, timed steps and end-to-end leg)."""
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        rows = [s for (t, s) in list(self.samples) if t0 <= t <= t1 + 0.06] or [s for (_, s) in list(self.samples)]
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
    (the same OpenCV functions; falls back to the oracle's C restatement if cv2 is missing), then - with a network size -
    cv::resize of both eyes, then the C/OpenMP restatement of the reference's preprocess and reproject kernels (which have no
    CPU code in the reference), tensors in the workload's dtype.  Returns (pairs, seconds, how)."""
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
    net = tuple(wl["net"]) if wl.get("net") else None
    tdt = np.float16 if wl["dtype"] == "f16" else np.float32
    keys = (("intrinsic_left", "distCoeffs_left", "R_L", "P1"), ("intrinsic_right", "distCoeffs_right", "R_R", "P2"))
    Q = calib["Q"] if net is None else calib["Q_net"]
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
        if net is not None:
            rect = [cv2.resize(x, net, interpolation=cv2.INTER_LINEAR) if cv2 else OC.resize(x, net) for x in rect]
        OC.preprocess_eye(mid, rect[0], tdt)
        OC.preprocess_eye(mid, rect[1], tdt)
        OC.reproject(mid, D[i % len(D)], rect[0], Q)
        done += 1
        el = time.perf_counter() - t0
        if (max_pairs and done >= max_pairs) or (budget_s and el >= budget_s):
            return done, el, ("cv2 " + cv2.__version__) if cv2 else "oracle_c"


def net_size(wl):
    return tuple(wl["net"]) if wl.get("net") else (wl["W"], wl["H"])


def workload_config(wname, wl, world, reproject="dense", heatmap=False):
    """The `config` object: identical for both arms (the driver compares them)."""
    w, h = net_size(wl)
    return {"workload": wname, "model": wl["model"], "size": [wl["W"], wl["H"]], "net_size": [w, h], "pairs_per_gpu_per_step": wl["batch"],
            "tensor_dtype": wl["dtype"], "reproject": reproject, "heatmap": bool(heatmap),
            "parallelism": f"frame-sharded x{world}, no collective",
            "l2": "working set per step %.0f MB >> 126 MB L2 (no flush needed)" % (bytes_per_pair(wl, "total") * wl["batch"] / 1e6)}


def with_net_Q(wl, calib):
    """Reprojection happens at network size: Q of the resized camera."""
    if wl.get("net"):
        from oracle import oracle as O
        sx = wl["net"][0] / wl["W"]
        calib = dict(calib)
        calib["Q_net"] = O.scale_calibration(calib, sx)["Q"]
    return calib


def run_reference_arm(args, wl, wname):
    """--impl reference: the reference's CPU path on the host cores, rank 0 only.  Same config as our arm: a step is the same
    number of pairs of the same workload with the same tensor type; the other ranks exit without work."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle as O
    threads = os.cpu_count() or 1
    calib = with_net_Q(wl, load_calib(wl))
    n = wl["batch"]
    w, h = net_size(wl)
    uniq = min(n, 16)                                         # distinct input frames cycled through a step (host memory, not timing)
    L, R = make_inputs(wl, uniq)
    D = [O.synthetic_disparity(w, h, 30 + i) * (-1 if wl["model"] == "RAFTStereo" else 1) for i in range(4)]
    # one step = n pairs, like our arm; bounded so that K + W steps end within a few minutes even at 1080p (about 25 pairs/s):
    # a step longer than ~6 s is cut to the pairs that fit and the rate extrapolated - said in `sample`.
    cpu_reference_pairs(wl, calib, L, R, D, threads, max_pairs=2)
    p0, s0, how = cpu_reference_pairs(wl, calib, L, R, D, threads, max_pairs=min(n, 8))
    per_step = max(1, min(n, int(6.0 * p0 / s0)))
    for _ in range(args.warmup):
        cpu_reference_pairs(wl, calib, L, R, D, threads, max_pairs=per_step)
    t = 0.0
    for _ in range(args.steps):
        _, el, how = cpu_reference_pairs(wl, calib, L, R, D, threads, max_pairs=per_step)
        t += el
    value = args.steps * per_step / t
    desc = f"{per_step} of the {n} pairs of a step of {wname}, {args.steps} steps, {threads} threads (batch-1 calls, maps recomputed every frame as the " \
           f"reference does; {how} for initUndistortRectifyMap+remap{'+resize' if wl.get('net') else ''}, oracle C/OpenMP for preprocess+reproject, " \
           f"{wl['dtype']} tensors)"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * (t / args.steps) * (n / per_step), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8 fixed-point + f32 (%s tensors)" % wl["dtype"], "data": "synthetic",
            "config": workload_config(wname, wl, int(os.environ.get("WORLD_SIZE", "1"))),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": desc},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------------------------
REF_SYMS = {   # mangled names of the reference's host launchers in oracle/_ref/libref_<Model>.so (oracle/build_ref.py)
    "RAFTStereo": ("_Z21RAFTStereo_preprocessPhPfiiP11CUstream_st", "_Z23cuda_reprojectImageTo3DPhPfS0_S0_ii"),
    "CREStereo": ("_Z20CREStereo_preprocessPhPfiiP11CUstream_st", "_Z28CREStereo_reprojectImageTo3DPhPfS0_S0_ii"),
    "FastACVNet_plus": ("_Z26FastACVNet_plus_preprocessPhPfiiP11CUstream_st", "_Z34FastACVNet_plus_reprojectImageTo3DPhPfS0_S0_ii"),
}


def time_reference_kernels(wl, n, reps=5):
    """the reference's OWN two CUDA kernels, compiled unmodified for sm_100a (oracle/_ref), on device-resident
    frames of this workload - n pairs = 2n preprocess launches + n reproject launches per step, batch 1 each, as the reference
    issues them (RAFTStereo/src/stereo_preprocess.cu:72-87).  Context for stage A's normalise half and for stage B; they have no
    rectification (that is CPU code in the reference) and f32 tensors only.  Not part of `value`."""
    import ctypes as C
    import torch
    from oracle import build_ref
    model = wl["model"]
    path = build_ref.lib_path(model)
    if model not in REF_SYMS or not os.path.exists(path):
        return None
    lib = C.CDLL(path)
    vp, ci = C.c_void_p, C.c_int
    pre, rep = getattr(lib, REF_SYMS[model][0]), getattr(lib, REF_SYMS[model][1])
    pre.argtypes = [vp, vp, ci, ci, vp]
    rep.argtypes = [vp, vp, vp, vp, ci, ci]
    w, h = net_size(wl)
    m = min(n, 16)                                            # distinct frames; the launch count per step is still 2n + n
    img = torch.randint(0, 256, (2, m, h, w, 3), dtype=torch.uint8, device="cuda")
    ten = torch.empty((2, m, 3, h, w), dtype=torch.float32, device="cuda")
    disp = torch.rand((m, h, w), dtype=torch.float32, device="cuda") * 60 + 1
    pts = torch.empty((m, h * w, 6), dtype=torch.float32, device="cuda")
    q = torch.eye(4, dtype=torch.float32, device="cuda").reshape(16).contiguous()
    torch.cuda.synchronize()
    cur = torch.cuda.current_stream()
    legacy = torch.cuda.ExternalStream(0)                     # the reference's reproject launcher uses the legacy default stream
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    pre_ms, rep_ms = [], []
    for r in range(reps + 1):
        ev[0].record(legacy)
        for i in range(n):
            for e in range(2):
                pre(img[e, i % m].data_ptr(), ten[e, i % m].data_ptr(), w, h, None)
        ev[1].record(legacy)
        for i in range(n):
            rep(img[0, i % m].data_ptr(), disp[i % m].data_ptr(), pts[i % m].data_ptr(), q.data_ptr(), h, w)
        ev[2].record(legacy)
        torch.cuda.synchronize()
        if r:
            pre_ms.append(ev[0].elapsed_time(ev[1]))
            rep_ms.append(ev[1].elapsed_time(ev[2]))
    del cur
    p, r = statistics.median(pre_ms), statistics.median(rep_ms)
    wh = w * h
    return {"preprocess_ms": p, "reproject_ms": r, "launches_per_step": 3 * n, "pairs_per_step": n,
            "preprocess_GBps": 2 * n * wh * 15 / (p * 1e-3) / 1e9, "reproject_GBps": n * wh * (31 + (4 if model == "RAFTStereo" else 0)) / (r * 1e-3) / 1e9,
            "what": f"oracle/_ref/libref_{model}.so = the reference's .cu compiled unmodified for sm_100a; f32 tensors, no rectification, batch-1 launches"}


def host_copy_ceiling(host_in, host_out, reps, barrier):
    """What the box can do for the SAME traffic with nothing but copies: per step, every input array from pinned memory on one
    stream and every output array into pinned memory on another, concurrently, bare cudaMemcpyAsync in blocks of <= 64 MB,
    all ranks at once.  host_in / host_out: flat uint8 torch views of the pinned arrays of the e2e leg.  Returns seconds per step
    (this rank)."""
    import torch
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    blk = 64 << 20
    dev_in = [torch.empty(t.numel(), dtype=torch.uint8, device="cuda") for t in host_in]
    dev_out = [torch.empty(t.numel(), dtype=torch.uint8, device="cuda") for t in host_out]

    def one():
        with torch.cuda.stream(s_in):
            for d, hbuf in zip(dev_in, host_in):
                for o in range(0, hbuf.numel(), blk):
                    d[o:o + blk].copy_(hbuf[o:o + blk], non_blocking=True)
        with torch.cuda.stream(s_out):
            for d, hbuf in zip(dev_out, host_out):
                for o in range(0, hbuf.numel(), blk):
                    hbuf[o:o + blk].copy_(d[o:o + blk], non_blocking=True)
    one()
    torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        one()
        s_in.synchronize()
        s_out.synchronize()
    return (time.perf_counter() - t0) / reps


def run_workload(args, wname, wl, rank, world, local, headline, sampler=None):
    """Times one workload on this rank's GPU; returns the result dict on rank 0 (None elsewhere)."""
    import torch
    import torch.distributed as dist
    from stereoalgorithms_b200 import _lib as SL
    from stereoalgorithms_b200 import sharding
    from stereoalgorithms_b200.geometry import COMPACT, DENSE, F16, F32, PinnedArray, StereoGeometry

    n = wl["batch"]
    if args.strong and world > 1:                            # fixed total batch, contiguous shard per rank (sa_shard_range's partition)
        from stereoalgorithms_b200.sharding import shard_range
        b_, e_ = shard_range(wl["batch"], rank, world)
        n = max(e_ - b_, 1)
    W, H = wl["W"], wl["H"]
    w, h = net_size(wl)
    npix = w * h
    calib = with_net_Q(wl, load_calib(wl))
    L, R = make_inputs(wl, n, rank)
    dL, dR = torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()
    in_pitch = 0
    if (3 * W) % 16 and args.pitched_rows:
        # device frames as cudaMallocPitch / cv::cuda::GpuMat lay them out: every row starts 16-byte aligned (1242-wide: 3726 -> 3728
        # bytes per row), which is what the tensor-copy kernel needs (with --opt bulk=2); default: packed rows
        in_pitch = (3 * W + 15) // 16 * 16
        def pitched(t):
            p = torch.zeros((n, H, in_pitch), dtype=torch.uint8, device="cuda")
            p[:, :, :3 * W] = t.reshape(n, H, 3 * W)
            return p
        dL, dR = pitched(dL), pitched(dR)
    compact = args.reproject == "compact"

    g = StereoGeometry(W, H, model=wl["model"], dtype=F16 if wl["dtype"] == "f16" else F32, max_batch=n, gpu_id=local, flags=SL.FLAG_TIMING,
                       net_size=wl.get("net"))
    g.set_calibration(calib)
    if wl.get("net"):
        g.set_Q(calib["Q_net"])
    if args.frame_fastest >= 0:
        g.set_option("frame_fastest", args.frame_fastest)
    for kv in args.opt:
        k_, v_ = kv.split("=")
        g.set_option(k_, int(v_))
    g.stub_disparity(n, seed=1234 + rank)
    g.synchronize()

    stream = torch.cuda.Stream()                              # a real (non-null) stream: NULL would mean "the handle's own stream"
    sp = stream.cuda_stream                                   # every kernel of the step is launched on this stream
    assert sp != 0
    counts = torch.zeros(n, dtype=torch.int32, device="cuda") if compact else None

    def step():
        g.rectify_preprocess(dL, dR, n=n, pitch=in_pitch, stream=sp)
        g.reproject(n, COMPACT if compact else DENSE, counts=counts, to_host=False, stream=sp)
        if args.heatmap:
            g.heatmap(n, to_host=False, stream=sp)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    t_workload0 = time.time()
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
    t_wall1 = time.time()
    step_time = ev0.elapsed_time(ev1) / 1e3                   # device time of exactly K steps, events on the launch stream
    launches = g.launch_count - launches0
    rect_ms, rect_cnt = g.stage_time(SL.STAGE_RECTIFY)
    resize_ms = 0.0
    if wl.get("net"):
        resize_ms = g.stage_time(SL.STAGE_RESIZE)[0]         # separate resize / preprocess launches, if the path is not fused
        rect_ms += resize_ms                                  # stage A = everything up to the tensors
    rep_ms, rep_cnt = g.stage_time(SL.STAGE_REPROJECT)
    heat_ms, heat_cnt = g.stage_time(SL.STAGE_HEATMAP) if args.heatmap else (0.0, 0)
    dev_ms = rect_ms + rep_ms + heat_ms                      # sum of the per-stage event pairs
    kept = float(npix)                                       # points kept per frame (compact mode)
    if compact:
        # RAFTStereo's kernel flips the disparity's sign in place, so consecutive steps see +d (94 % of the pixels valid) and -d (1 %):
        # the timed loop is the average of the two, and so is the byte count
        c1 = float(counts.sum().item())
        step()
        torch.cuda.synchronize()
        kept = (c1 + float(counts.sum().item())) / 2 / n
    pairs_total, step_time, launches = sharding.aggregate(n * args.steps, step_time, launches)     # SUM pairs, MAX time, SUM launches
    _, dev_s, _ = sharding.aggregate(0, dev_ms / 1e3)
    value = pairs_total / step_time

    def bpp(stage):
        b = bytes_per_pair(wl, stage)
        if compact and stage in ("reproject", "total", "total_heatmap"):
            b -= int(24 * (npix - kept))                      # only the kept rows are written
        return b

    # ---- e2e: host buffers through the reference-facing call --------------------------------------------------------
    e2e = None
    if not args.no_e2e:
        k2 = args.e2e_steps or min(args.steps, 20)
        hin = [PinnedArray(L.shape, np.uint8), PinnedArray(R.shape, np.uint8)]
        hin[0].array[:] = L
        hin[1].array[:] = R
        hout = [PinnedArray((n, npix, 6), np.float32), PinnedArray((n, h, w), np.float32), PinnedArray(L.shape, np.uint8), PinnedArray(R.shape, np.uint8)]
        hheat = PinnedArray((n, h, w, 3), np.uint8) if args.heatmap else None

        def e2e_step():
            g.run_batch(hin[0].array, hin[1].array, n, rectify=True, stub_seed=0, pointcloud=hout[0].array, disparity=hout[1].array,
                        rect_left=hout[2].array, rect_right=hout[3].array, heatmap=hheat.array if hheat else None)

        for _ in range(3):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(k2):
            e2e_step()                                        # synchronous: returns when the results are in host memory
        el2 = time.perf_counter() - t0
        pairs2, t2, _ = sharding.aggregate(n * k2, el2)
        h2d = 2 * n * W * H * 3
        d2h = n * (npix * (24 + 4) + W * H * 6 + (npix * 3 if args.heatmap else 0))
        e2e = {"value": pairs2 / t2, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": k2,
               "api": "sa_run_batch (pinned host images in; point cloud + disparity + rectified L/R%s out)" % (" + heat map" if args.heatmap else ""),
               "pcie_gbs_per_gpu": (h2d + d2h) * k2 / t2 / 1e9}
        # the same bytes per step with nothing but copies, all ranks at once: what this host can do for this traffic
        try:
            flat = lambda arr: torch.from_numpy(arr.array.reshape(-1).view(np.uint8))   # pinned already (cudaMallocHost): plain views
            sec = host_copy_ceiling([flat(a) for a in hin], [flat(a) for a in hout + ([hheat] if hheat else [])], max(3, k2 // 2), barrier)
            _, sec, _ = sharding.aggregate(0, sec)
            ceil_pairs = n * world / sec
            e2e["host_copy_ceiling"] = {"value": ceil_pairs, "unit": UNIT, "gbs_per_gpu": (h2d + d2h) / sec / 1e9,
                                        "what": "bare pinned cudaMemcpyAsync of the same bytes per step (H2D and D2H on two streams, 64 MB blocks), all ranks at once"}
            e2e["frac_of_host_ceiling"] = e2e["value"] / ceil_pairs
            torch.cuda.empty_cache()
        except Exception as ex:                               # the ceiling is context, never a reason to lose the bench line
            e2e["host_copy_ceiling"] = {"error": repr(ex)[:200]}
        for a in hin + hout + ([hheat] if hheat else []):
            a.free()

    res = None
    if rank == 0:
        peak, peak_src = measured_hbm_peak()
        stages = {"rectify": (rect_ms, rect_cnt), "reproject": (rep_ms, rep_cnt)}
        if args.heatmap:
            stages["heatmap"] = (heat_ms, heat_cnt)
        dom = max(("rectify", "reproject"), key=lambda k: stages[k][0])
        dom_ms, dom_cnt = stages[dom]
        per_launch_bytes = bpp(dom) * n
        achieved = per_launch_bytes / (dom_ms / dom_cnt * 1e-3) / 1e9
        total_key = "total_heatmap" if args.heatmap else "total"
        kernel_names = {"rectify": "rectify_ws_kernel + resize_norm_kernel (or rectify_ws_kernel<RESIZE> with --opt fused_resize=1)" if wl.get("net")
                        else "rectify_ws_kernel / rectify_batch_kernel (fused rectify+preprocess)",
                        "reproject": "reproject_compact_kernel (+ count, scan)" if compact else "reproject_dense_kernel"}
        roof = {"bound": "hbm", "kernel": kernel_names[dom],
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src, "traffic": None,
                "algorithmic_bytes_per_launch": per_launch_bytes, "avg_launch_ms": dom_ms / dom_cnt,
                "stages": {k: {"ms_per_launch": v[0] / max(v[1], 1), "GBps": bpp(k) * n / (v[0] / max(v[1], 1) * 1e-3) / 1e9,
                               "frac": bpp(k) * n / (v[0] / max(v[1], 1) * 1e-3) / 1e9 / peak, "bytes_per_pair": bpp(k)} for k, v in stages.items()},
                "resize_preprocess_ms_per_step": resize_ms / max(rect_cnt, 1) if wl.get("net") else None,
                "whole_step": {"GBps": bpp(total_key) * n * args.steps / step_time / 1e9,
                               "frac": bpp(total_key) * n * args.steps / step_time / 1e9 / peak, "bytes_per_pair": bpp(total_key)}}
        prof = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(prof):
            try:
                roof["traffic"] = json.load(open(prof)).get(wname, {}).get(dom)
            except Exception:
                pass
        res = {"value": value, "unit": UNIT, "ms_per_step": 1e3 * step_time / args.steps, "kernel_ms_per_step": 1e3 * dev_s / args.steps,
               "host_ms_per_step": 1e3 * elapsed / args.steps, "gpu_launches": launches, "roofline": roof, "e2e": e2e,
               "config": dict(workload_config(wname, wl, world, args.reproject, args.heatmap), **({"device_row_pitch": in_pitch} if in_pitch else {})),
               "_wall": (t_workload0, time.time()),
               "dtype": "u8 fixed-point + f32 (%s tensors)" % wl["dtype"]}
        if compact:
            res["points_kept_per_frame"] = kept
        if headline and not args.no_ref_kernels and world == 1:
            try:
                res["reference_kernels"] = time_reference_kernels(wl, n)
            except Exception as ex:
                res["reference_kernels"] = {"error": repr(ex)[:200]}
    g.close()
    del dL, dR
    torch.cuda.empty_cache()
    return res, (calib, L, R)


# ---------------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=list(WORKLOADS), help="time only this workload (default: the headline + EXTRA_WORKLOADS)")
    ap.add_argument("--batch", type=int, default=0, help="override the per-GPU batch of the workload")
    ap.add_argument("--e2e-steps", type=int, default=0, help="steps of the host-to-host leg (default min(steps, 20))")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="headline workload only")
    ap.add_argument("--no-ref-kernels", action="store_true", help="skip the reference-kernels-on-B200 row")
    ap.add_argument("--heatmap", action="store_true", help="include stage C in the step")
    ap.add_argument("--strong", action="store_true", help="strong scaling: the workload's batch is the TOTAL, sharded frame-wise over the ranks (default: per GPU)")
    ap.add_argument("--reproject", default="dense", choices=["dense", "compact"], help="stage B mode (compact = validity filter + compaction, extension a9)")
    ap.add_argument("--pitched-rows", action="store_true", help="device frames whose rows are not a multiple of 16 bytes (1242 wide) with a 16-byte-aligned row pitch, not packed")
    ap.add_argument("--frame-fastest", type=int, default=-1)
    ap.add_argument("--opt", action="append", default=[], help="library tuning option key=value (sa_set_option), repeatable")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    wname = args.workload or DEFAULT_WORKLOAD
    extras = [] if (args.workload or args.no_extra or args.batch) else list(EXTRA_WORKLOADS)
    wl = dict(WORKLOADS[wname])
    if args.batch:
        wl["batch"] = args.batch

    if args.impl == "reference":
        run_reference_arm(args, wl, wname)
        return

    import torch
    import torch.distributed as dist

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

    sampler = ClockSampler(local) if rank == 0 else None
    head, (calib, L, R) = run_workload(args, wname, wl, rank, world, local, headline=True)
    others = {}
    for name in extras:
        r, _ = run_workload(args, name, dict(WORKLOADS[name]), rank, world, local, headline=False)
        if r is not None:
            r["clocks"] = sampler.window(*r.pop("_wall")) if sampler else None
            others[name] = r

    if rank == 0:
        clocks = sampler.window(*head.pop("_wall")) if sampler else None      # the headline workload's own window, not the extras'
        if sampler:
            sampler.stop()
        n, W, H = wl["batch"], wl["W"], wl["H"]
        w, h = net_size(wl)
        cpu = None
        if not args.no_cpu_baseline and world == 1:
            from oracle import oracle as O
            threads = os.cpu_count() or 1
            m = min(n, 16)
            D = [O.synthetic_disparity(w, h, 30 + i) * (-1 if wl["model"] == "RAFTStereo" else 1) for i in range(4)]
            cpu_reference_pairs(wl, calib, L[:m], R[:m], D, threads, max_pairs=4)
            pairs, secs, how = cpu_reference_pairs(wl, calib, L[:m], R[:m], D, threads, budget_s=12.0)
            pairs_c, secs_c, _ = cpu_reference_pairs(wl, calib, L[:m], R[:m], D, threads, budget_s=4.0, cached_maps=True)
            cpu = {"value": pairs / secs, "unit": UNIT, "cores": threads, "kind": "port",
                   "sample": f"{pairs} pairs of {wname} in {secs:.1f} s, batch-1 calls, maps recomputed every frame as the reference does "
                             f"({how} for initUndistortRectifyMap+remap, oracle C/OpenMP for preprocess+reproject, {wl['dtype']} tensors)",
                   "maps_cached_value": pairs_c / secs_c}
        cfg = head.pop("config")
        notes = {"numa_node_rank0": numa,
                 "network_slot": "synthetic disparity written once before timing (RAFTStereo: its sign alternates per step, same work)"}
        line = {"metric": METRIC, "value": head.pop("value"), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": head.pop("ms_per_step"), "higher_is_better": True, "scaling": "strong" if args.strong else "weak", "vs_baseline": None,
                "dtype": head.pop("dtype"), "data": "synthetic", "config": cfg}
        line.update(head)
        line.update({"cpu_baseline": cpu, "clocks": clocks, "notes": notes})
        if others:
            line["workloads"] = others
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
