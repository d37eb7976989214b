"""Feature extraction with the reference's call signatures, computed by the sm_100a kernel.

Mirrors ``kws/libs/audio_processor.py`` of the reference:
  * ``AudioConfig``                       (``:20-72``)   -- same fields and defaults
  * ``AudioProcessor.extract_features``   (``:235-278``) -- same keyword arguments, same
    ``x or config`` defaulting, same ``nfft = max(config.fft_size, int(winlen*samplerate))``
    override of the ``nfft`` argument, then the psf 0.6 ``mfcc`` defaults (rectangular window,
    preemph 0.97, ceplifter 22, appendEnergy, lowfreq 0, highfreq sr/2)
  * ``AudioProcessor.to_mfcc``            (``kws/libs/signal_handler.py:112-131``, legacy twin)

Additions: a leading batch dimension (``[B, N] -> [B, nframes, C]``, each row equal to the
per-clip result), CUDA tensors in / CUDA float32 tensors out, int16 PCM input, and
``extract_logfbank`` (psf ``logfbank``).  numpy in -> numpy out (float64 dtype like psf, values
computed in float32 on the device).  File loading, time shift and noise mixing
(``:103-233``) are host-side augmentation and stay out of scope.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from pathlib import Path
from typing import Any, Dict, Optional

import numpy as np
import torch

from . import _native
from .common.errors import AudioProcessingError


@dataclass
class AudioConfig:
    """Same record as the reference ``AudioConfig`` (audio_processor.py:20-72)."""

    background_volume: float = 0.1
    background_frequency: float = 0.8
    use_background_noise: bool = True
    sample_rate: int = 16000
    clip_duration_ms: int = 1000
    time_shift_ms: float = 100
    frame_length: float = 0.025
    frame_step: float = 0.01
    num_cepstral_coeffs: int = 10
    num_mel_filters: int = 26
    fft_size: int = 512

    @property
    def time_shift(self) -> int:
        return int((self.time_shift_ms * self.sample_rate) / 1000)

    @property
    def desired_samples(self) -> int:
        return int(self.sample_rate * self.clip_duration_ms / 1000)

    def to_dict(self) -> Dict[str, Any]:
        return {k: getattr(self, k) for k in (
            "background_volume", "background_frequency", "use_background_noise", "sample_rate",
            "clip_duration_ms", "time_shift_ms", "frame_length", "frame_step",
            "num_cepstral_coeffs", "num_mel_filters", "fft_size")}


class _FeaturePlan:
    """One native handle configured for one (sr, frame_len, hop, nfft, nfilt, numcep) tuple."""

    def __init__(self, device_index: int, key, preemph: Optional[float] = None):
        lib = _native.lib()
        self.handle = _native.Handle(device_index)
        cfg = _native.FeatureConfig()
        lib.kws_feature_config_default(C.byref(cfg))
        (cfg.sample_rate, cfg.frame_len, cfg.frame_step, cfg.nfft, cfg.nfilt, cfg.numcep) = key
        if preemph is not None:
            cfg.preemph = preemph
        self.preemph = float(cfg.preemph)
        _native.check(lib.kws_features_configure(self.handle.ptr, C.byref(cfg)), AudioProcessingError)
        self.frame_len, self.frame_step, self.nfilt = int(cfg.frame_len), int(cfg.frame_step), int(cfg.nfilt)
        self.ncep = int(lib.kws_features_ncep(self.handle.ptr))


class AudioProcessor:
    """Drop-in for the feature half of the reference ``AudioProcessor``."""

    def __init__(self, dataset_path: Optional[Path] = None, config: Optional[AudioConfig] = None,
                 device: Optional[torch.device] = None):
        self.config = config if config is not None else AudioConfig()
        self.dataset_path = dataset_path
        self.device = torch.device(device) if device is not None else torch.device("cuda", 0)
        if self.device.type != "cuda":
            raise AudioProcessingError("kws_b200 runs on CUDA devices only (no CPU fallback)")
        self._plans: Dict[tuple, _FeaturePlan] = {}

    # -- plumbing -----------------------------------------------------------------------
    def _dev_index(self) -> int:
        return self.device.index if self.device.index is not None else torch.cuda.current_device()

    def _plan(self, samplerate, numcep, winlen, winstep, nfilt) -> _FeaturePlan:
        lib = _native.lib()
        nfft = max(self.config.fft_size, int(winlen * samplerate))      # audio_processor.py:268
        key = (int(samplerate), int(lib.kws_round_half_up(winlen * samplerate)),
               int(lib.kws_round_half_up(winstep * samplerate)), int(nfft), int(nfilt), int(numcep))
        plan = self._plans.get(key)
        if plan is None:
            plan = _FeaturePlan(self._dev_index(), key)
            self._plans[key] = plan
        return plan

    def _run(self, signal, plan: _FeaturePlan, logmel: bool):
        lib = _native.lib()
        is_numpy = not isinstance(signal, torch.Tensor)
        x = torch.as_tensor(np.asarray(signal)) if is_numpy else signal
        if x.dim() not in (1, 2):
            raise AudioProcessingError(f"expected a [N] or [B, N] signal, got shape {tuple(x.shape)}")
        single = x.dim() == 1
        if single:
            x = x.unsqueeze(0)
        if x.shape[1] < 1:
            raise AudioProcessingError("empty signal")
        i16 = x.dtype == torch.int16
        if not i16:
            x = x.to(torch.float32)
        x = x.to(self.device, non_blocking=True).contiguous()
        B, N = x.shape
        nframes = int(lib.kws_num_frames(N, plan.frame_len, plan.frame_step))
        cols = plan.nfilt if logmel else plan.ncep
        out = torch.empty((B, nframes, cols), dtype=torch.float32, device=self.device)
        if B > 0:
            stream = _native.current_stream_ptr(self._dev_index())
            if logmel:
                fn = lib.kws_logfbank_i16 if i16 else lib.kws_logfbank
            else:
                fn = lib.kws_mfcc_i16 if i16 else lib.kws_mfcc
            with torch.cuda.device(self.device):
                _native.check(fn(plan.handle.ptr, C.c_void_p(x.data_ptr()), B, N, x.stride(0),
                                 C.c_void_p(out.data_ptr()), stream), AudioProcessingError)
        if single:
            out = out[0]
        if is_numpy:
            return out.cpu().numpy().astype(np.float64)
        return out

    def sliding_features_incremental(self, recording: torch.Tensor, hop_samples: int,
                                     window_samples: Optional[int] = None, **kw) -> torch.Tensor:
        """Same result as :meth:`sliding_features` (``[nwin, nframes, numcep]``) with the frames consecutive windows
        share computed ONCE: the recording is transformed as one long clip, the two frames per
        window that differ from the recording's frame grid (first: no pre-emphasis predecessor; last: zero tail past
        the window end) are transformed separately, and the windows are assembled by a gather.  Needs
        ``hop_samples`` to be a multiple of the frame step.  Frames 1..T-2 equal the per-window call bit for bit;
        the two edge frames come from the generic kernel and agree within fp32 rounding (north_star tolerance)."""
        lib = _native.lib()
        sr = kw.get("samplerate") or self.config.sample_rate
        numcep = kw.get("numcep") or self.config.num_cepstral_coeffs
        winlen = kw.get("winlen") or self.config.frame_length
        winstep = kw.get("winstep") or self.config.frame_step
        nfilt = kw.get("nfilt") or self.config.num_mel_filters
        plan = self._plan(sr, numcep, winlen, winstep, nfilt)
        fl, fs = plan.frame_len, plan.frame_step
        N, hop = int(window_samples or self.config.desired_samples), int(hop_samples)
        x = recording.to(self.device, torch.float32).contiguous()
        if x.dim() != 1 or x.shape[0] < N or hop < 1 or N < fl:
            raise AudioProcessingError(f"expected a [L] recording with L >= {N} >= frame length and hop >= 1")
        if hop % fs != 0:
            raise AudioProcessingError(f"incremental features need hop_samples to be a multiple of the frame step ({fs})")
        L = x.shape[0]
        nwin = (L - N) // hop + 1
        T = int(lib.kws_num_frames(N, fl, fs))
        if T < 3:
            return self.sliding_features(recording, hop_samples, window_samples, **kw)
        nfft = max(self.config.fft_size, int(winlen * sr))
        ekey = (int(sr), fl, fl, int(nfft), int(nfilt), int(numcep), 0.0)     # one frame per row, no pre-emphasis
        eplan = self._plans.get(ekey)
        if eplan is None:
            eplan = self._plans[ekey] = _FeaturePlan(self._dev_index(), ekey[:6], preemph=0.0)
        stream = _native.current_stream_ptr(self._dev_index())
        # the recording is only needed up to the last window's end
        Lr = (nwin - 1) * hop + N
        rec_frames = int(lib.kws_num_frames(Lr, fl, fs))
        rec_feats = torch.empty((rec_frames, plan.ncep), dtype=torch.float32, device=self.device)
        edge_pcm = torch.empty((2 * nwin, fl), dtype=torch.float32, device=self.device)
        edge_feats = torch.empty((2 * nwin, plan.ncep), dtype=torch.float32, device=self.device)
        out = torch.empty((nwin, T, plan.ncep), dtype=torch.float32, device=self.device)
        vp = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
        with torch.cuda.device(self.device):
            chk = lambda rc: _native.check(rc, AudioProcessingError)  # noqa: E731
            chk(lib.kws_mfcc(plan.handle.ptr, vp(x), 1, Lr, Lr, vp(rec_feats), stream))
            chk(lib.kws_stream_edge_frames(plan.handle.ptr, vp(x), L, nwin, hop, N, fl, fs, plan.preemph, vp(edge_pcm), stream))
            chk(lib.kws_mfcc(eplan.handle.ptr, vp(edge_pcm), 1, 2 * nwin * fl, 2 * nwin * fl, vp(edge_feats), stream))
            chk(lib.kws_stream_gather_windows(plan.handle.ptr, vp(rec_feats), rec_frames, vp(edge_feats), nwin, T, plan.ncep,
                                              hop // fs, vp(out), stream))
        return out

    def sliding_features(self, recording: torch.Tensor, hop_samples: int, window_samples: Optional[int] = None,
                         **kw) -> torch.Tensor:
        """MFCCs of every ``window_samples`` window taken every ``hop_samples`` from a device-resident
        recording ``[L]``: ``[nwin, nframes, numcep]``, row ``k`` = ``extract_features(recording[k*hop:...])``.
        The windows are read in place as overlapping rows."""
        lib = _native.lib()
        plan = self._plan(kw.get("samplerate") or self.config.sample_rate,
                          kw.get("numcep") or self.config.num_cepstral_coeffs,
                          kw.get("winlen") or self.config.frame_length,
                          kw.get("winstep") or self.config.frame_step,
                          kw.get("nfilt") or self.config.num_mel_filters)
        N, hop = int(window_samples or self.config.desired_samples), int(hop_samples)
        x = recording.to(self.device, torch.float32).contiguous()
        if x.dim() != 1 or x.shape[0] < N or hop < 1:
            raise AudioProcessingError(f"expected a [L] recording with L >= {N} and hop >= 1")
        nwin = (x.shape[0] - N) // hop + 1
        nframes = int(lib.kws_num_frames(N, plan.frame_len, plan.frame_step))
        out = torch.empty((nwin, nframes, plan.ncep), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            _native.check(lib.kws_mfcc(plan.handle.ptr, C.c_void_p(x.data_ptr()), nwin, N, hop,
                                       C.c_void_p(out.data_ptr()), _native.current_stream_ptr(self._dev_index())),
                          AudioProcessingError)
        return out

    # -- reference API --------------------------------------------------------------------
    def extract_features(self, signal, samplerate: int = None, numcep: int = None, nfft: int = None,
                         winlen: float = None, winstep: float = None, nfilt: int = None):
        """MFCC features, ``[nframes, min(numcep, nfilt)]`` (or ``[B, nframes, .]``)."""
        samplerate = samplerate or self.config.sample_rate
        numcep = numcep or self.config.num_cepstral_coeffs
        nfft = nfft or self.config.fft_size      # accepted and ignored, like the reference (:268)
        winlen = winlen or self.config.frame_length
        winstep = winstep or self.config.frame_step
        nfilt = nfilt or self.config.num_mel_filters
        return self._run(signal, self._plan(samplerate, numcep, winlen, winstep, nfilt), logmel=False)

    def extract_logfbank(self, signal, samplerate: int = None, winlen: float = None,
                         winstep: float = None, nfilt: int = None):
        """psf ``logfbank`` ("log-mel"): the intermediate of ``mfcc`` just before the DCT."""
        samplerate = samplerate or self.config.sample_rate
        winlen = winlen or self.config.frame_length
        winstep = winstep or self.config.frame_step
        nfilt = nfilt or self.config.num_mel_filters
        return self._run(signal, self._plan(samplerate, nfilt, winlen, winstep, nfilt), logmel=True)

    def to_mfcc(self, signal, samplerate=16000, numcep=10, nfft=512, winlen=0.025, winstep=0.01,
                nfilt=26):
        """Legacy spelling (signal_handler.py:112-131): fixed defaults, ``max(512, .)`` override."""
        lib = _native.lib()
        key = (int(samplerate), int(lib.kws_round_half_up(winlen * samplerate)),
               int(lib.kws_round_half_up(winstep * samplerate)), max(512, int(winlen * samplerate)),
               int(nfilt), int(numcep))
        plan = self._plans.get(key)
        if plan is None:
            plan = self._plans[key] = _FeaturePlan(self._dev_index(), key)
        return self._run(signal, plan, logmel=False)

    def extract_features_host(self, pcm: torch.Tensor, out: Optional[torch.Tensor] = None, **kw):
        """End-to-end variant on HOST tensors (ideally pinned): chunked H2D -> kernel -> D2H
        pipelined inside the library (``kws_mfcc_host``).  ``pcm`` float32 ``[B, N]`` on the CPU."""
        lib = _native.lib()
        if pcm.device.type != "cpu" or pcm.dtype not in (torch.float32, torch.int16) or pcm.dim() != 2:
            raise AudioProcessingError("extract_features_host expects a float32 (or int16 PCM) [B, N] CPU tensor")
        pcm = pcm.contiguous()
        plan = self._plan(kw.get("samplerate") or self.config.sample_rate,
                          kw.get("numcep") or self.config.num_cepstral_coeffs,
                          kw.get("winlen") or self.config.frame_length,
                          kw.get("winstep") or self.config.frame_step,
                          kw.get("nfilt") or self.config.num_mel_filters)
        B, N = pcm.shape
        nframes = int(lib.kws_num_frames(N, plan.frame_len, plan.frame_step))
        if out is None:
            out = torch.empty((B, nframes, plan.ncep), dtype=torch.float32, pin_memory=True)
        with torch.cuda.device(self.device):
            fn = lib.kws_mfcc_host_i16 if pcm.dtype == torch.int16 else lib.kws_mfcc_host
            _native.check(fn(plan.handle.ptr, C.c_void_p(pcm.data_ptr()), B, N, C.c_void_p(out.data_ptr())),
                          AudioProcessingError)
        return out

    # -- augmentation: the stages of the reference ``transform`` before the features ----
    def set_background_data(self, clips) -> None:
        """Device-resident noise bank = the reference's ``background_data`` list (audio_processor.py:103-128),
        concatenated; clips shorter than ``desired_samples`` are read modulo their length (``np.tile``)."""
        clips = [torch.as_tensor(c, dtype=torch.float32).reshape(-1) for c in clips]
        if any(c.numel() == 0 for c in clips):
            raise AudioProcessingError("empty background clip")
        self._bank_len = torch.tensor([c.numel() for c in clips], dtype=torch.int64)
        self._bank_start = torch.cumsum(self._bank_len, 0) - self._bank_len
        self._bank = torch.cat(clips).to(self.device) if clips else None

    def draw_augmentation(self, labels: torch.Tensor, generator: Optional[torch.Generator] = None) -> Dict[str, torch.Tensor]:
        """Per-clip random draws with the reference's distributions (audio_processor.py:181-183, 207-231):
        shift ~ U{-time_shift .. time_shift-1}; noise clip uniform; offset ~ U{0 .. len'-N-1} over the tiled
        length; volume ~ U(0,1) for silence, else U(0, background_volume) with probability
        background_frequency, else 0.  (The reference uses the global numpy/random state; parity is
        distributional)"""
        cfg, N = self.config, self.config.desired_samples
        labels = torch.as_tensor(labels).to("cpu", torch.int64)
        B = labels.numel()
        ts = cfg.time_shift
        g = generator
        shift = torch.randint(-ts, ts, (B,), generator=g, dtype=torch.int32) if ts > 0 else torch.zeros(B, dtype=torch.int32)
        silence = (labels == 0)
        draws = {"shift": shift, "silence": silence.to(torch.uint8)}
        bank = getattr(self, "_bank", None)
        if bank is not None:
            idx = torch.randint(0, self._bank_len.numel(), (B,), generator=g)
            ln = self._bank_len[idx]
            tiled = torch.where(ln <= N, ((N + ln - 1) // ln) * ln, ln)
            span = torch.clamp(tiled - N, min=1)
            off = (torch.rand(B, generator=g, dtype=torch.float64) * span).to(torch.int64).clamp_(max=span - 1)
            u1, u2 = torch.rand(B, generator=g), torch.rand(B, generator=g)
            vol = torch.where(silence, u2, torch.where(u1 < cfg.background_frequency, u2 * cfg.background_volume,
                                                       torch.zeros(B)))
            use = silence | torch.tensor(bool(cfg.use_background_noise))
            draws.update({"noise_start": self._bank_start[idx], "noise_len": torch.where(use, ln, torch.zeros_like(ln)),
                          "noise_offset": off, "noise_volume": vol.float()})
        return draws

    def augment(self, pcm: torch.Tensor, draws: Dict[str, torch.Tensor]) -> torch.Tensor:
        """Apply silence zeroing, time shift and background-noise mix on the device (``kws_augment``);
        ``pcm[B, N]`` float32 CUDA, ``draws`` from :meth:`draw_augmentation` (or hand-made)."""
        lib = _native.lib()
        x = pcm.to(self.device, torch.float32)
        if x.dim() != 2 or x.stride(1) != 1:
            x = x.reshape(-1, x.shape[-1]).contiguous()
        B, N = x.shape
        dev = {k: v.to(self.device).contiguous() for k, v in draws.items()}
        out = torch.empty((B, N), dtype=torch.float32, device=self.device)
        bank = getattr(self, "_bank", None) if "noise_start" in dev else None
        ptr = lambda k: C.c_void_p(dev[k].data_ptr()) if k in dev else None  # noqa: E731
        plan = self._plan(self.config.sample_rate, self.config.num_cepstral_coeffs, self.config.frame_length,
                          self.config.frame_step, self.config.num_mel_filters)
        with torch.cuda.device(self.device):
            _native.check(lib.kws_augment(
                plan.handle.ptr, C.c_void_p(x.data_ptr()), B, N, x.stride(0), ptr("shift"), ptr("silence"),
                C.c_void_p(bank.data_ptr()) if bank is not None else None, ptr("noise_start"), ptr("noise_len"),
                ptr("noise_offset"), ptr("noise_volume"), C.c_void_p(out.data_ptr()),
                _native.current_stream_ptr(self._dev_index())), AudioProcessingError)
        return out

    def launch_count(self) -> int:
        return sum(p.handle.launch_count() for p in self._plans.values())
