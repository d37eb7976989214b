"""BASELINE config 4: log-mel + CRNN/GRU inference, batch 16384: clips/s of PCM -> log-mel -> CRNN logits.
usage: bench_crnn.py [B] [reps]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import kws_b200  # noqa: F401,E402
from kws_b200.audio_processor import AudioConfig, AudioProcessor  # noqa: E402
from kws_b200.models import CRNN  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
torch.manual_seed(0)
m = CRNN(n_features=40).cuda()
proc = AudioProcessor(None, AudioConfig(), device="cuda:0")
pcm = torch.rand(B, 16000, device="cuda") * 2 - 1


def timed(fn):
    for _ in range(2):
        out = fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out


ms_f, feats = timed(lambda: proc.extract_logfbank(pcm, nfilt=40))
x = feats.unsqueeze(1)
ms_m, _ = timed(lambda: m(x))
ms_all, _ = timed(lambda: m(proc.extract_logfbank(pcm, nfilt=40).unsqueeze(1)))
print(json.dumps({"workload": "logmel40 + CRNN(conv 16ch, GRU 64, 1 layer), batch %d" % B,
                  "logmel_clips_per_s": B / ms_f * 1e3, "crnn_clips_per_s": B / ms_m * 1e3,
                  "pcm_to_logits_clips_per_s": B / ms_all * 1e3, "ms": {"logmel": ms_f, "crnn": ms_m, "both": ms_all}}))
