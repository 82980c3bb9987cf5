"""The N > 1 host logic on CPU: world_size 2 over gloo (the data path itself has no collective)."""
import os
import socket

import pytest

from stereoalgorithms_b200.sharding import shard_range


def test_shard_range_partitions_everything():
    for n in (0, 1, 7, 64, 256, 513):
        for world in (1, 2, 3, 4, 8):
            seen = []
            for r in range(world):
                b, e = shard_range(n, r, world)
                assert 0 <= b <= e <= n
                seen += list(range(b, e))
            assert seen == list(range(n))                       # contiguous, ordered, nothing lost or duplicated
            sizes = [shard_range(n, r, world)[1] - shard_range(n, r, world)[0] for r in range(world)]
            assert max(sizes) == -(-n // world)
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def test_library_partition_is_the_same_partition():
    """sa_shard_range (what sa_run_batch_multi cuts a host batch with) and sharding.shard_range (what bench.py's ranks use)."""
    import ctypes as C
    from stereoalgorithms_b200 import _lib
    lib = _lib.load()
    b, e = C.c_int(), C.c_int()
    for n in (0, 1, 7, 64, 256, 512, 513, 1024):
        for world in (1, 2, 3, 4, 8):
            for r in range(world):
                assert lib.sa_shard_range(n, r, world, C.byref(b), C.byref(e)) == 0
                assert (b.value, e.value) == shard_range(n, r, world)
    assert lib.sa_shard_range(10, 2, 2, C.byref(b), C.byref(e)) == _lib.SA_ERR_INVALID_ARG
    assert lib.sa_shard_range(10, 0, 0, C.byref(b), C.byref(e)) == _lib.SA_ERR_INVALID_ARG


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import torch.distributed as dist
    from stereoalgorithms_b200.sharding import aggregate, shard_range, throughput
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    b, e = shard_range(101, rank, world)
    secs = 0.5 if rank == 0 else 2.0                            # rank 1 is the slow one
    dist.barrier()
    pairs, tmax, launches = aggregate(e - b, secs, launches_local=2 * (e - b))
    thr = throughput(e - b, secs)
    q.put((rank, b, e, pairs, tmax, launches, thr))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_gloo_aggregate_max_time_sum_pairs():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, b0, e0, pairs0, t0, l0, thr0), (r1, b1, e1, pairs1, t1, l1, thr1) = out
    assert (b0, e0, b1, e1) == (0, 51, 51, 101)
    assert pairs0 == pairs1 == 101 and t0 == t1 == 2.0 and l0 == l1 == 202       # SUM of pairs, MAX of time, on every rank
    assert thr0 == thr1 == pytest.approx(101 / 2.0)
