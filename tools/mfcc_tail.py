"""How much of the north_star MFCC tolerance (1e-4 + 1e-3 |ref|) the fp32 kernel uses against the float64
oracle, over many clips: quantiles of |gpu - oracle| / bound.  usage: mfcc_tail.py [clips] [kind]"""
import os
import sys
from concurrent.futures import ProcessPoolExecutor

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import kws_b200  # noqa: F401,E402
from kws_b200.audio_processor import AudioConfig, AudioProcessor  # noqa: E402
from kws_b200.utils import synth  # noqa: E402
from oracle import psf_mfcc  # noqa: E402


def _oracle(chunk):
    return psf_mfcc.extract_features_batch(chunk, numcep=40, nfilt=40)


if __name__ == "__main__":
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
    kind = sys.argv[2] if len(sys.argv) > 2 else "uniform"
    if kind == "uniform":
        pcm = synth.uniform_pcm(B, seed=11)
    else:  # speech-like: 1/f-shaped noise bursts, the spectrum pre-emphasis is meant for
        g = torch.Generator().manual_seed(11)
        w = torch.randn(B, 16000, generator=g)
        spec = torch.fft.rfft(w) / torch.arange(1, 8002).float().sqrt()
        pcm = torch.fft.irfft(spec, n=16000)
        pcm = (pcm / pcm.abs().amax(dim=1, keepdim=True) * 0.5).contiguous()
    proc = AudioProcessor(None, AudioConfig(), device="cuda:0")
    got = proc.extract_features(pcm.cuda(), numcep=40, nfilt=40).cpu().numpy().astype(np.float64)
    chunks = np.array_split(pcm.numpy(), max(1, os.cpu_count() or 1) * 4)
    with ProcessPoolExecutor() as ex:
        want = np.concatenate(list(ex.map(_oracle, chunks)))
    use = np.abs(got - want) / (1e-4 + 1e-3 * np.abs(want))
    q = np.quantile(use, [0.5, 0.99, 0.9999, 0.999999, 1.0])
    print({"clips": B, "kind": kind, "values": int(use.size), "max_abs_err": float(np.abs(got - want).max()),
           "tolerance_use": {"p50": q[0], "p99": q[1], "p99.99": q[2], "p99.9999": q[3], "max": q[4]},
           "outside": int((use > 1).sum())})
