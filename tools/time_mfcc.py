"""CUDA-event timing of the fused MFCC kernel alone: clips/s at B clips (float32 and int16 PCM)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import kws_b200  # noqa: F401,E402
from kws_b200.audio_processor import AudioConfig, AudioProcessor  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
proc = AudioProcessor(None, AudioConfig(), device="cuda:0")
pcm = torch.rand((B, 16000), device="cuda") * 2 - 1
i16 = (pcm * 32767).round().to(torch.int16)
for name, x, kw in (("f32 40x40", pcm, dict(numcep=40, nfilt=40)), ("i16 40x40", i16, dict(numcep=40, nfilt=40)),
                    ("f32 10x26", pcm, dict())):
    for _ in range(3):
        proc.extract_features(x, **kw)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    torch.cuda.synchronize()
    ev[0].record()
    for _ in range(reps):
        proc.extract_features(x, **kw)
    ev[1].record()
    torch.cuda.synchronize()
    ms = ev[0].elapsed_time(ev[1]) / reps
    print(f"{name}: {ms:.3f} ms per {B} clips = {B / ms / 1e3:.3f} M clips/s")
