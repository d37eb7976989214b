"""Timing of the model forward (features resident in HBM) and of PCM -> logits: clips/s of the fused
tcgen05 kernel vs the layered fp32 kernels.  usage: bench_model.py [B] [F] [reps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import kws_b200  # noqa: F401,E402
from kws_b200.audio_processor import AudioConfig  # noqa: E402
from kws_b200.models import DepthwiseSeparableConv  # noqa: E402
from kws_b200.pipeline import KeywordSpotter  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
F = int(sys.argv[2]) if len(sys.argv) > 2 else 40
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
paths = sys.argv[4].split(",") if len(sys.argv) > 4 else ["auto", "layered"]
torch.manual_seed(0)
m = DepthwiseSeparableConv().cuda()
with torch.no_grad():
    for p in m.parameters():
        if p.dim() == 1:
            p.normal_(0, 0.1)
x = torch.randn(B, 1, 99, F, device="cuda") * 4


def timed(fn):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out


res = {}
for path in paths:
    m.set_kernel_path(path)
    ms, out = timed(lambda: m(x))
    res[path] = out
    print(f"model forward [{path:7s}] B={B} F={F}: {ms:.3f} ms  {B / ms * 1e3 / 1e6:.3f} M clips/s")
if len(res) == 2:
    a, b = res["auto"], res["layered"]
    print("fused vs layered: max abs diff %.3e, max |logit| %.3f, argmax equal %s" % (
        (a - b).abs().max().item(), b.abs().max().item(), bool((a.argmax(1) == b.argmax(1)).all())))
m.set_kernel_path("auto")
if F == 40:
    ks = KeywordSpotter(m, AudioConfig(), numcep=40, nfilt=40, device="cuda:0")
    pcm = torch.rand(B, 16000, device="cuda") * 2 - 1
    ms, _ = timed(lambda: ks(pcm))
    print(f"pcm -> logits (mfcc kernel + fused model kernel) B={B}: {ms:.3f} ms  {B / ms * 1e3 / 1e6:.3f} M clips/s")
