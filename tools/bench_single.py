"""Single-launch PCM -> logits vs the two kernels back to back, and a sweep of the role split.
usage: bench_single.py [B]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import kws_b200  # noqa: F401,E402
from kws_b200.audio_processor import AudioConfig  # noqa: E402
from kws_b200.models import DepthwiseSeparableConv  # noqa: E402
from kws_b200.pipeline import KeywordSpotter  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
torch.manual_seed(0)
m = DepthwiseSeparableConv().cuda()
with torch.no_grad():
    for p in m.parameters():
        if p.dim() == 1:
            p.normal_(0, 0.1)
pcm = torch.rand(B, 16000, device="cuda") * 2 - 1


def timed(fn, reps=5):
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


for numcep, nfilt in ((40, 40), (10, 26)):
    ks = KeywordSpotter(m, AudioConfig(), numcep=numcep, nfilt=nfilt, device="cuda:0")
    ks.configure_single_launch(min_clips=0)
    ms = timed(lambda: ks(pcm))
    print(f"F={numcep} B={B} two kernels: {ms:.3f} ms {B / ms / 1e3:.3f} M clips/s")
    for chunk in (8, 16, 32, 64, 128):
        ks.configure_single_launch(min_clips=1, chunk_clips=chunk)
        ms = timed(lambda: ks(pcm))
        print(f"F={numcep} B={B} single launch chunk={chunk}: {ms:.3f} ms {B / ms / 1e3:.3f} M clips/s")
for Bs in (256, 1024, 4096, 16384):
    ks = KeywordSpotter(m, AudioConfig(), numcep=40, nfilt=40, device="cuda:0")
    x = pcm[:Bs]
    ks.configure_single_launch(min_clips=0)
    a = timed(lambda: ks(x), 30)
    ks.configure_single_launch(min_clips=1, chunk_clips=32)
    b = timed(lambda: ks(x), 30)
    print(f"B={Bs}: two kernels {a * 1e3:.1f} us, single launch {b * 1e3:.1f} us")
