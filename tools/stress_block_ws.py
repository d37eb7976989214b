"""Stress of the warp-specialised block kernel (model_gemm.cu): many forward passes of the same batch must be
bit-identical (a race in the A ring, the weight ring or the accumulator-slot ring shows up as a differing run), and
the result must match the layered fp32 kernels.  usage: stress_block_ws.py [runs]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import kws_b200  # noqa: F401
from kws_b200.models import DepthwiseSeparableConv
runs = int(sys.argv[1]) if len(sys.argv) > 1 else 20
bad = 0
for channels, blocks, T, F, B in [(276, 5, 99, 40, 777), (172, 4, 99, 40, 333), (64, 3, 99, 40, 1500), (256, 2, 61, 64, 211)]:
    torch.manual_seed(channels)
    m = DepthwiseSeparableConv(channels=channels, num_blocks=blocks).cuda()
    x = torch.randn(B, 1, T, F, device="cuda") * 2.0
    ref = m(x).clone()
    diff = 0
    for _ in range(runs):
        y = m(x)
        diff += int(not torch.equal(y, ref))
    m.set_kernel_path("layered")
    lay = m(x)
    err = (ref - lay).abs().max().item()
    print(f"{channels}x{blocks} B={B}: {diff} of {runs} runs differ, max |fast - layered| = {err:.2e}")
    bad += diff + int(err > 2e-3)
print("OK" if bad == 0 else "FAILED")
sys.exit(1 if bad else 0)
