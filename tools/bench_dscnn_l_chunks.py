"""DS-CNN-L forward: one call of 8192 clips (four passes of 2048 through the workspace inside the C call) against four
calls of 2048 clips on different / on the same slice.  usage: bench_dscnn_l_chunks.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import kws_b200  # noqa: F401
from kws_b200.models import DepthwiseSeparableConv
torch.manual_seed(0)
m = DepthwiseSeparableConv(channels=276, num_blocks=5).cuda()
x = torch.randn(8192, 1, 99, 40, device="cuda")
parts = [x[i * 2048:(i + 1) * 2048].contiguous() for i in range(4)]


def timed(fn, reps=3):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


print("one call of 8192      : %.2f ms" % timed(lambda: m(x)))
print("four calls, 4 slices  : %.2f ms" % timed(lambda: [m(p) for p in parts]))
print("four calls, same slice: %.2f ms" % timed(lambda: [m(parts[0]) for _ in range(4)]))
