"""One PCM -> logits call (single launch) for profiling: run_single.py [B] [chunk]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import kws_b200  # noqa: F401
from kws_b200.audio_processor import AudioConfig
from kws_b200.models import DepthwiseSeparableConv
from kws_b200.pipeline import KeywordSpotter
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4736
torch.manual_seed(0)
m = DepthwiseSeparableConv().cuda()
ks = KeywordSpotter(m, AudioConfig(), numcep=40, nfilt=40, device="cuda:0")
ks.configure_single_launch(min_clips=1, chunk_clips=int(sys.argv[2]) if len(sys.argv) > 2 else 32)
pcm = torch.rand(B, 16000, device="cuda") * 2 - 1
for _ in range(3):
    out = ks(pcm)
torch.cuda.synchronize()
print(out.shape)
