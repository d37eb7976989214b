"""Host -> device copy ceiling of the box, for 1..N concurrent ranks (one process per GPU under torchrun):
what the end-to-end (host buffer) path of bench.py can reach at most.  Every rank copies a pinned buffer to its
own GPU at the same time (barrier before, max time over ranks); reported: GB/s per rank and aggregate, for plain
pinned and write-combined pinned source buffers, with and without a concurrent device -> host stream.

    python tools/pcie_probe.py                       # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/pcie_probe.py
"""
import ctypes
import json
import os
import time

import torch
import torch.distributed as dist

rank, local, world = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("LOCAL_RANK", 0), ("WORLD_SIZE", 1)))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)

n = 1 << 30
d = torch.empty(n, dtype=torch.uint8, device=dev)
d2 = torch.empty(n // 4, dtype=torch.uint8, device=dev)
h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
h.fill_(1)
h2 = torch.empty(n // 4, dtype=torch.uint8, pin_memory=True)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def sync():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def maxr(x):
    t = torch.tensor([x], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def timed(fn, reps=5):
    fn()
    sync()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return maxr(time.perf_counter() - t0) / reps


def h2d():
    with torch.cuda.stream(s1):
        d.copy_(h, non_blocking=True)


def both():
    with torch.cuda.stream(s1):
        d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2):
        h2.copy_(d2, non_blocking=True)


res = {"world": world, "bytes_per_copy": n}
t = timed(h2d)
res["h2d_gbs_per_rank"] = n / t / 1e9
res["h2d_gbs_aggregate"] = world * n / t / 1e9
t = timed(both)
res["h2d_with_d2h_gbs_aggregate"] = world * n / t / 1e9

# write-combined pinned source (cudaHostAllocWriteCombined = 4): the CPU never reads it back, DMA reads skip the snoop
try:
    rt = ctypes.CDLL("libcudart.so.12")
    ptr = ctypes.c_void_p()
    assert rt.cudaHostAlloc(ctypes.byref(ptr), ctypes.c_size_t(n), ctypes.c_uint(4)) == 0
    ctypes.memset(ptr, 1, n)
    st = torch.cuda.current_stream().cuda_stream

    def h2d_wc():
        assert rt.cudaMemcpyAsync(ctypes.c_void_p(d.data_ptr()), ptr, ctypes.c_size_t(n), ctypes.c_int(1), ctypes.c_void_p(st)) == 0

    t = timed(h2d_wc)
    res["h2d_write_combined_gbs_aggregate"] = world * n / t / 1e9
    rt.cudaFreeHost(ptr)
except Exception as e:  # pragma: no cover
    res["h2d_write_combined_gbs_aggregate"] = repr(e)

try:
    res["cpu_affinity"] = len(os.sched_getaffinity(0))
    res["numa_nodes"] = len([x for x in os.listdir("/sys/devices/system/node") if x.startswith("node")])
except Exception:
    pass
if rank == 0:
    print(json.dumps(res))
if world > 1:
    dist.destroy_process_group()
