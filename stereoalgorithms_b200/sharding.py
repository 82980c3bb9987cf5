"""Frame-wise sharding of a batch of stereo pairs over the GPUs of one box.

Frames share no data, so there is no collective on the data path: every rank (one process per GPU) takes a contiguous
block of pairs and keeps its own copy of the per-calibration constants.  torch.distributed is used only to agree on the
clock: a barrier around the timed region, the MAX of the per-rank elapsed times and the SUM of the pairs processed.
"""
from __future__ import annotations


def shard_range(n_total: int, rank: int, world: int) -> tuple[int, int]:
    """[begin, end) of the pairs rank `rank` owns: contiguous blocks of ceil(n_total / world), the last ones shorter
    (possibly empty).  Same partition as the library's sa_shard_range / sa_run_batch_multi (tests/test_multi_rank_cpu.py)."""
    if world <= 0 or not (0 <= rank < world) or n_total < 0:
        raise ValueError(f"bad shard request: n_total={n_total} rank={rank} world={world}")
    per = -(-n_total // world)
    b = min(n_total, rank * per)
    return b, min(n_total, b + per)


def aggregate(pairs_local: int, seconds_local: float, launches_local: int = 0, device=None):
    """-> (pairs_total, seconds_max, launches_total) over all ranks; identity when torch.distributed is not initialised."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return pairs_local, seconds_local, launches_local
    dev = device if device is not None else ("cuda" if dist.get_backend() == "nccl" else "cpu")
    t = torch.tensor([float(seconds_local)], dtype=torch.float64, device=dev)
    s = torch.tensor([float(pairs_local), float(launches_local)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(s, op=dist.ReduceOp.SUM)
    return int(round(float(s[0]))), float(t[0]), int(round(float(s[1])))


def throughput(pairs_local: int, seconds_local: float, device=None) -> float:
    """Whole-job pairs/s: all pairs of all ranks divided by the slowest rank's time."""
    pairs, secs, _ = aggregate(pairs_local, seconds_local, 0, device)
    return pairs / secs if secs > 0 else 0.0
