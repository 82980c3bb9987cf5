"""bench.py's bookkeeping (no GPU): the algorithmic bytes it divides by are the SURVEY.md section 8d figures, the workload table names
the BASELINE configs, and the command line the driver uses parses."""
import importlib.util
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench():
    spec = importlib.util.spec_from_file_location("bench_module", os.path.join(ROOT, "bench.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_algorithmic_bytes_match_the_scope_table():
    b = _bench()
    wh = 640 * 480
    raft, fast = b.WORKLOADS["raft_realtime_640x480_b64_f16"], b.WORKLOADS["fastacv_640x480_b64_f32"]
    assert b.bytes_per_pair(fast, "rectify") == 36 * wh and b.bytes_per_pair(raft, "rectify") == 24 * wh        # stage A: f32 / f16 tensors
    assert b.bytes_per_pair(fast, "reproject") == 31 * wh and b.bytes_per_pair(raft, "reproject") == 35 * wh    # stage B (+ RAFT write-back)
    assert b.bytes_per_pair(fast, "total") == 67 * wh and b.bytes_per_pair(raft, "total") == 59 * wh
    rz = b.WORKLOADS["stream_1080p_resize_640x480_b64_f32"]
    WH = 1920 * 1080
    assert b.bytes_per_pair(rz, "rectify") == 12 * WH + 24 * wh + 3 * wh   # raw in + rectified out at full size, tensors + the left u8 at network size
    assert b.bytes_per_pair(rz, "reproject") == 31 * wh               # disparity / point cloud live at network size


def test_heatmap_step_counts_stage_c_and_both_arms_describe_the_same_config():
    b = _bench()
    wh = 640 * 480
    raft = b.WORKLOADS["raft_realtime_640x480_b64_f16"]
    assert b.bytes_per_pair(raft, "heatmap") == 7 * wh and b.bytes_per_pair(raft, "total_heatmap") == 66 * wh
    cfg = b.workload_config("raft_realtime_640x480_b64_f16", raft, 1)
    assert cfg["workload"] == "raft_realtime_640x480_b64_f16" and cfg["pairs_per_gpu_per_step"] == 64 and cfg["tensor_dtype"] == "f16"
    for name in b.EXTRA_WORKLOADS:                       # the default run also times the 1080p half of the metric
        assert name in b.WORKLOADS
    assert "stream_1080p_b64_f32" in b.EXTRA_WORKLOADS


def test_reference_arm_runs_the_same_config_on_the_cpu():
    """--impl reference on a tiny batch: same config object as our arm, bounded, CPU only."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1", "--batch", "2"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    b = _bench()
    wl = dict(b.WORKLOADS[b.DEFAULT_WORKLOAD], batch=2)
    assert line["impl"] == "reference" and line["unit"] == "pairs/s" and line["value"] > 0
    assert line["config"] == b.workload_config(b.DEFAULT_WORKLOAD, wl, 1)
    assert line["cpu_baseline"]["kind"] == "port" and line["e2e"]["h2d_bytes_per_step"] == 0 and line["gpu_launches"] == 0


def test_default_workload_is_the_config_the_metric_is_quoted_on():
    b = _bench()
    base = json.load(open(os.path.join(ROOT, "BASELINE.json")))
    assert b.DEFAULT_WORKLOAD == "raft_realtime_640x480_b64_f16"
    wl = b.WORKLOADS[b.DEFAULT_WORKLOAD]
    assert (wl["W"], wl["H"], wl["batch"], wl["dtype"], wl["model"]) == (640, 480, 64, "f16", "RAFTStereo")
    assert b.UNIT == "pairs/s" and isinstance(base, dict)
    for name in ("fastacv_kitti_1242x375_b256_f32", "stream_1080p_b64_f32", "uhd_2160p_b16_f32"):
        assert name in b.WORKLOADS


def test_command_line_the_driver_uses_parses():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--help"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0
    for flag in ("--gpus", "--steps", "--warmup", "--impl"):
        assert flag in r.stdout
