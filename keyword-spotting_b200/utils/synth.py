"""Seeded synthetic inputs.  Pure numpy/torch-CPU; no CUDA, no oracle."""

import numpy as np
import torch

CLIP_SAMPLES = 16000


def uniform_pcm(batch: int, n: int = CLIP_SAMPLES, seed: int = 0) -> torch.Tensor:
    """``U(-1,1)`` float32 ``[batch, n]`` from a CPU generator so the CPU oracle and the
    GPU see identical bits."""
    g = torch.Generator().manual_seed(seed)
    return torch.rand(batch, n, generator=g, dtype=torch.float32) * 2.0 - 1.0


def parity_pcm(batch: int = 64, n: int = CLIP_SAMPLES, seed: int = 0) -> torch.Tensor:
    """The parity set: uniform clips plus the special clips SURVEY.md 8(d) lists."""
    assert batch >= 8
    x = uniform_pcm(batch, n, seed)
    t = torch.arange(n, dtype=torch.float64)
    x[0] = 0.0                                             # all-zero clip (silence label)
    x[1] = 0.0
    x[1, 0] = 0.75                                         # zero except sample 0
    x[2] = (0.5 * torch.sin(2 * np.pi * 1000.0 * t / 16000.0)).float()   # 1 kHz sine
    g = torch.Generator().manual_seed(seed + 1)
    x[3] = torch.randn(n, generator=g) * 0.1               # N(0, 0.1)
    x[4] = torch.round(x[4] * 32767.0) / 32768.0           # int16-quantised (librosa.load)
    rs = np.random.RandomState(0)                          # reference test's signal
    x[5] = torch.from_numpy(rs.rand(n).astype(np.float32))
    x[6] = 1.0                                             # DC
    x[7, n // 2:] = 0.0                                    # half clip silent
    return x
