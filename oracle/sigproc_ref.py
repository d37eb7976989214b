"""sigproc oracle: numpy float64 restatement of the reference's torch helpers.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

Restates ``/root/reference/kws/libs/speech_features/sigproc.py``:
``framesig`` (``:14-52``), ``magspec`` (``:55-77``), ``powspec`` (``:80-90``),
``preemphasis`` (``:93-103``).  Differences from the psf functions of the same
name (``oracle/psf_mfcc.py``): ``frame_len``/``frame_step`` are plain ints (no
``round_half_up``), the default window is the *periodic* Hann window with its
values created in float32 (``torch.hann_window(frame_len)``), and ``magspec``
crops explicitly when the frame is longer than NFFT.

Pinned by ``tests/golden/sigproc_seed0.npz`` which ``tests/golden/make_golden.py``
produces by running the reference module itself on the reference test's input
(``/root/reference/tests/kws/libs/speech_features/test_sigproc.py:8-21``).
"""

from __future__ import annotations

import math

import numpy as np


def hann_periodic_f32(n: int) -> np.ndarray:
    """``torch.hann_window(n)`` (periodic=True, float32 values)."""
    # torch builds the window in float32: arange * (2*pi/n) -> cos -> * -0.5 + 0.5.  The
    # cancellation near the ends makes this differ from a float64 window by up to 2.4e-7
    # absolute; restated in float32 it matches torch to 1 ulp of 0.5 (6e-8).  GPU tests pass
    # the window values explicitly to both sides, so this only matters for the golden check.
    k = np.arange(n, dtype=np.float32)
    c = np.cos(k * np.float32(2.0 * np.pi / n)).astype(np.float32)
    return (c * np.float32(-0.5) + np.float32(0.5)).astype(np.float32)


def num_frames(signal_len: int, frame_len: int, frame_step: int) -> int:
    if signal_len <= frame_len:
        return 1
    return 1 + math.ceil((signal_len - frame_len) / frame_step)


def framesig(signal: np.ndarray, frame_len: int, frame_step: int, window=None) -> np.ndarray:
    """``window=None`` means the reference default (periodic Hann, fp32 values);
    pass an explicit array (e.g. ones) for any other ``winfunc``."""
    n = signal.shape[0]
    nf = num_frames(n, frame_len, frame_step)
    pad_len = (nf - 1) * frame_step + frame_len
    pad = np.zeros(pad_len, dtype=signal.dtype)
    pad[:n] = signal
    idx = np.arange(frame_len)[None, :] + (np.arange(nf) * frame_step)[:, None]
    frames = pad[idx]
    win = hann_periodic_f32(frame_len) if window is None else np.asarray(window)
    return frames * win


def magspec(frames: np.ndarray, nfft: int) -> np.ndarray:
    if frames.shape[1] > nfft:
        frames = frames[:, :nfft]
    return np.abs(np.fft.rfft(frames, n=nfft, axis=1))


def powspec(frames: np.ndarray, nfft: int) -> np.ndarray:
    return 1.0 / nfft * np.square(magspec(frames, nfft))


def preemphasis(signal: np.ndarray, coeff: float = 0.97) -> np.ndarray:
    return np.concatenate([signal[:1], signal[1:] - coeff * signal[:-1]])
