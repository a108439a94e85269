"""Multi-GPU layout on CPU: world_size-2 gloo processes shard the streams, gather per-stream records
and reduce the box-wide strongest peak exactly like bench.py does over NCCL (no data-path collective)."""
import importlib
import os
import sys

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, total, q):
    sys.path.insert(0, ROOT)
    sdrb = importlib.import_module("sdr-for-android-lib_b200")
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    first, count = sdrb.shard_streams(total, world, rank)
    rec = np.zeros((count, 3), dtype=sdrb.RESULT_DTYPE)
    for s in range(count):
        for b in range(3):
            rec[s, b]["absPeakDb"] = ((first + s) * 7919 + b * 104729) % 1000 / 10.0
            rec[s, b]["peakBin"] = first + s
    allrec = sdrb.gather_records(rec, dist)
    peak = sdrb.strongest_peak(rec, first, dist)
    q.put((rank, first, count, allrec["peakBin"][:, 0].tolist(), peak))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("total", [7, 8])
def test_two_rank_sharding(total):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000) + total
    procs = [ctx.Process(target=_worker, args=(r, 2, port, total, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, f0, c0, all0, pk0), (r1, f1, c1, all1, pk1) = out
    assert f0 == 0 and f1 == c0 and c0 + c1 == total and abs(c0 - c1) <= 1
    assert all0 == all1 == list(range(total))             # gathered in global stream order on every rank
    exp = max((((s * 7919 + b * 104729) % 1000) / 10.0, -s) for s in range(total) for b in range(3))
    assert pk0 == pk1 and abs(pk0[0] - exp[0]) < 1e-4 and pk0[1] == -exp[1]


def test_shard_streams_edges():
    sys.path.insert(0, ROOT)
    sdrb = importlib.import_module("sdr-for-android-lib_b200")
    assert [sdrb.shard_streams(10, 4, r) for r in range(4)] == [(0, 3), (3, 3), (6, 2), (8, 2)]
    assert sdrb.shard_streams(1, 8, 7) == (1, 0)
    with pytest.raises(ValueError):
        sdrb.shard_streams(4, 2, 2)
