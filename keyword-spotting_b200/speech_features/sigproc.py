"""CUDA drop-ins for the reference's torch signal-processing helpers.

Same names, arguments and results as ``kws/libs/speech_features/sigproc.py`` of the reference
(``framesig`` ``:14-52``, ``magspec`` ``:55-77``, ``powspec`` ``:80-90``, ``preemphasis``
``:93-103``), executed by the kernels in ``csrc/sigproc.cu``.  Tensors must live on a CUDA
device (float32 or float64, the dtype is followed like the torch code does); an optional
leading batch dimension is accepted.  There is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native
from ..common.errors import AudioProcessingError

_DTYPES = {torch.float32: 0, torch.float64: 1}


def _prep(t: torch.Tensor, what: str):
    if not isinstance(t, torch.Tensor):
        raise AudioProcessingError(f"{what}: expected a torch.Tensor")
    if t.device.type != "cuda":
        raise AudioProcessingError(f"{what}: tensor is on {t.device}; kws_b200 has no CPU fallback")
    if t.dtype not in _DTYPES:
        raise AudioProcessingError(f"{what}: dtype {t.dtype} not supported (float32/float64)")
    return t.contiguous(), _DTYPES[t.dtype]


def _stream(t):
    return _native.current_stream_ptr(t.device.index)


def preemphasis(signal: torch.Tensor, coeff: float = 0.97) -> torch.Tensor:
    """``cat(x[:1], x[1:] - coeff * x[:-1])`` along the last dimension."""
    x, dt = _prep(signal, "preemphasis")
    if x.dim() not in (1, 2):
        raise AudioProcessingError("preemphasis: expected [N] or [B, N]")
    out = torch.empty_like(x)
    B = 1 if x.dim() == 1 else x.shape[0]
    N = x.shape[-1]
    with torch.cuda.device(x.device):
        _native.check(_native.lib().kws_sigproc_preemphasis(
            dt, C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()), B, N, float(coeff), _stream(x)),
            AudioProcessingError)
    return out


def framesig(signal: torch.Tensor, frame_len: int, frame_step: int,
             winfunc: callable = torch.hann_window) -> torch.Tensor:
    """Frames ``[num_frames, frame_len]`` (``[B, num_frames, frame_len]`` for a batch): zero-padded
    to ``(num_frames-1)*frame_step + frame_len`` samples and multiplied by
    ``winfunc(frame_len, device=...)`` (default: periodic Hann created in float32)."""
    x, dt = _prep(signal, "framesig")
    if x.dim() not in (1, 2):
        raise AudioProcessingError("framesig: expected [N] or [B, N]")
    frame_len, frame_step = int(frame_len), int(frame_step)
    if frame_len < 1 or frame_step < 1:
        raise AudioProcessingError("framesig: frame_len and frame_step must be >= 1")
    lib = _native.lib()
    B = 1 if x.dim() == 1 else x.shape[0]
    N = x.shape[-1]
    nf = int(lib.kws_num_frames(N, frame_len, frame_step))
    win = winfunc(frame_len, device=x.device)
    out_dtype = torch.result_type(x, win)          # frames * win promotion, like the reference
    if out_dtype != x.dtype:
        x = x.to(out_dtype)
        dt = _DTYPES[out_dtype]
    win = win.to(out_dtype).contiguous()
    out = torch.empty((B, nf, frame_len), dtype=out_dtype, device=x.device)
    with torch.cuda.device(x.device):
        _native.check(lib.kws_sigproc_framesig(
            dt, C.c_void_p(x.data_ptr()), B, N, frame_len, frame_step, C.c_void_p(win.data_ptr()),
            C.c_void_p(out.data_ptr()), _stream(x)), AudioProcessingError)
    return out[0] if signal.dim() == 1 else out


def _spec(frames: torch.Tensor, NFFT: int, power: int, what: str) -> torch.Tensor:
    x, dt = _prep(frames, what)
    if x.dim() < 2:
        raise AudioProcessingError(f"{what}: expected [..., num_frames, frame_len]")
    NFFT = int(NFFT)
    if NFFT < 1:
        raise AudioProcessingError(f"{what}: NFFT must be >= 1")
    L = x.shape[-1]
    R = x.numel() // L if L else 0
    out = torch.empty(x.shape[:-1] + (NFFT // 2 + 1,), dtype=x.dtype, device=x.device)
    with torch.cuda.device(x.device):
        _native.check(_native.lib().kws_sigproc_spec(
            dt, C.c_void_p(x.data_ptr()), R, L, NFFT, power, C.c_void_p(out.data_ptr()), _stream(x)),
            AudioProcessingError)
    return out


def magspec(frames: torch.Tensor, NFFT: int) -> torch.Tensor:
    """``|rfft(frames, n=NFFT)|`` with frames cropped or right-zero-padded to NFFT."""
    return _spec(frames, NFFT, 0, "magspec")


def powspec(frames: torch.Tensor, NFFT: int) -> torch.Tensor:
    """``1/NFFT * magspec**2``."""
    return _spec(frames, NFFT, 1, "powspec")
