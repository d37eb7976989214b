"""World-size-2 gloo test (CPU) of the multi-rank host logic: shard ranges cover the batch once,
per-rank results reassemble in order, and timing is the MAX over ranks."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from kws_b200.utils.sharding import job_throughput, max_over_ranks, shard_range


def test_shard_range_properties():
    for total in (0, 1, 7, 256, 65536, 65537):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)
    assert job_throughput([100, 100], [1.0, 2.0]) == 100.0


def _worker(rank, world, port, total, out_q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard_range(total, rank, world)
    data = torch.arange(total, dtype=torch.float32)
    mine = data[lo:hi] * 2.0                          # stand-in for this rank's per-clip results
    sizes = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(sizes, torch.tensor([hi - lo]))
    bufs = [torch.zeros(int(s.item())) for s in sizes]
    # gather of results happens outside the timed region in bench.py; emulate with padded gather
    pad = max(int(s.item()) for s in sizes)
    padded = [torch.zeros(pad) for _ in range(world)]
    dist.all_gather(padded, torch.nn.functional.pad(mine, (0, pad - mine.numel())))
    whole = torch.cat([p[: b.numel()] for p, b in zip(padded, bufs)])
    t = max_over_ranks(1.0 + rank)                    # rank 1 is "slower"
    dist.barrier()
    if rank == 0:
        out_q.put((bool(torch.equal(whole, data * 2.0)), t))
    dist.destroy_process_group()


def test_two_rank_gloo():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, 1001, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok, t = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok and t == 2.0
