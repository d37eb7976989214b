"""Small driver for ncu: a few launches of the fused MFCC kernel on B synthetic clips."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import kws_b200  # noqa: F401,E402
from kws_b200.audio_processor import AudioConfig, AudioProcessor  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
proc = AudioProcessor(None, AudioConfig(), device="cuda:0")
pcm = torch.rand((B, 16000), device="cuda") * 2 - 1
for _ in range(reps):
    out = proc.extract_features(pcm, numcep=40, nfilt=40)
torch.cuda.synchronize()
print("ok", tuple(out.shape), float(out[0, 0, 0]))
