"""PCM -> logits in one call: the reference's ``extract_features`` (audio_processor.py:235-278)
followed by ``DepthwiseSeparableConv.forward`` (models.py:160-183), with the features never
leaving the device (``kws_pcm_to_logits``).  Batch-sharded across ranks by the caller: clips are
independent, so there is no collective on this path."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _native
from .audio_processor import AudioConfig
from .common.errors import AudioProcessingError, ModelError
from .models import DepthwiseSeparableConv


class KeywordSpotter:
    def __init__(self, model: DepthwiseSeparableConv, config: Optional[AudioConfig] = None,
                 numcep: Optional[int] = None, nfilt: Optional[int] = None,
                 device: Optional[torch.device] = None):
        self.model = model
        self.config = config if config is not None else AudioConfig()
        self.device = torch.device(device) if device is not None else torch.device("cuda", 0)
        if self.device.type != "cuda":
            raise ModelError("kws_b200 runs on CUDA devices only (no CPU fallback)")
        idx = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self._idx = idx
        lib = _native.lib()
        self._h = model.native_handle(self.device)          # model weights live on this handle
        cfg = _native.FeatureConfig()
        lib.kws_feature_config_default(C.byref(cfg))
        sr = self.config.sample_rate
        cfg.sample_rate = sr
        cfg.frame_len = int(lib.kws_round_half_up(self.config.frame_length * sr))
        cfg.frame_step = int(lib.kws_round_half_up(self.config.frame_step * sr))
        cfg.nfft = max(self.config.fft_size, int(self.config.frame_length * sr))
        cfg.nfilt = int(nfilt or self.config.num_mel_filters)
        cfg.numcep = int(numcep or self.config.num_cepstral_coeffs)
        self._cfg = cfg
        self._bind()

    def _bind(self) -> None:
        """The feature plan lives on the model's native handle: (re)install this spotter's configuration when
        another spotter of the same model, or a re-pack of the weights, was there last."""
        self._h = self.model.native_handle(self.device)     # re-pack if the weights changed
        if getattr(self.model, "_feature_cfg_owner", None) is not self or getattr(self, "_bound_h", None) is not self._h:
            _native.check(_native.lib().kws_features_configure(self._h.ptr, C.byref(self._cfg)), AudioProcessingError)
            self.model._feature_cfg_owner = self
            self._bound_h = self._h

    def configure_single_launch(self, min_clips: int = 1, chunk_clips: int = 0) -> None:
        """PCM -> logits runs as ONE kernel launch (``csrc/pcm_logits_fused.cu``: every CTA alternates between the
        MFCC of its next ``chunk_clips`` clips and the network on them, features in an L2-resident scratch) for
        batches of at least ``min_clips``; ``min_clips < 1`` always runs the MFCC kernel and the network kernel back
        to back.  ``chunk_clips = 0`` keeps the current chunk size (default 32)."""
        self._bind()
        _native.check(_native.lib().kws_pcm_to_logits_configure(self._h.ptr, int(min_clips), int(chunk_clips)), ModelError)

    def __call__(self, pcm: torch.Tensor) -> torch.Tensor:
        """``pcm[B, N]`` float32 (or int16 PCM, scaled by 1/32768 like ``librosa.load``) -> logits
        ``[B, num_classes]`` on the device."""
        lib = _native.lib()
        self._bind()
        i16 = pcm.dtype == torch.int16
        x = pcm.to(self.device, torch.int16 if i16 else torch.float32).contiguous()
        if x.dim() == 1:
            x = x.unsqueeze(0)
        B, N = x.shape
        out = torch.empty((B, self.model.num_classes), dtype=torch.float32, device=self.device)
        if B:
            with torch.cuda.device(self.device):
                _native.check((lib.kws_pcm_to_logits_i16 if i16 else lib.kws_pcm_to_logits)(
                    self._h.ptr, C.c_void_p(x.data_ptr()), B, N, x.stride(0), C.c_void_p(out.data_ptr()),
                    _native.current_stream_ptr(self._idx)), ModelError)
        return out

    def sliding_logits_incremental(self, recording: torch.Tensor, hop_samples: int,
                                   window_samples: Optional[int] = None) -> torch.Tensor:
        """:meth:`sliding_logits` with incremental features: the frames overlapping windows share
        are computed once (``AudioProcessor.sliding_features_incremental``; ``hop_samples`` must be a multiple of the
        frame step), then every window's feature map goes through the network.  ``recording`` is ``[L]``."""
        from .audio_processor import AudioProcessor
        if getattr(self, "_proc", None) is None:
            self._proc = AudioProcessor(None, self.config, device=self.device)
        feats = self._proc.sliding_features_incremental(recording, hop_samples, window_samples,
                                                        numcep=int(self._cfg.numcep), nfilt=int(self._cfg.nfilt))
        return self.model(feats.unsqueeze(1))

    def sliding_logits(self, recording: torch.Tensor, hop_samples: int,
                       window_samples: Optional[int] = None) -> torch.Tensor:
        """Sliding-window inference over long recordings (the batch form of the reference's
        1 s window concept, kws/inference/inference_local.py:114-188): every window of ``window_samples``
        (default ``config.desired_samples``) starting at ``k * hop_samples`` is classified.

        ``recording`` is ``[L]`` or ``[S, L]`` float32; returns ``[nwin, K]`` (``[S, nwin, K]``) with
        ``nwin = (L - window) // hop + 1``.  The windows are never materialised: the kernels read them in
        place as overlapping rows (row stride = hop), so HBM holds the recording once.  Row ``k`` equals
        ``self(recording[k*hop : k*hop + window])`` (bit for bit when hop is a multiple of 4 samples, i.e.
        when both run on the TMA-fed kernel)."""
        lib = _native.lib()
        self._bind()
        N = int(window_samples or self.config.desired_samples)
        hop = int(hop_samples)
        if hop < 1:
            raise AudioProcessingError("hop_samples must be >= 1")
        x = recording.to(self.device, torch.float32).contiguous()
        single = x.dim() == 1
        if single:
            x = x.unsqueeze(0)
        if x.dim() != 2 or x.shape[1] < N:
            raise AudioProcessingError(f"expected a [L] or [S, L] recording with L >= {N}, got {tuple(recording.shape)}")
        S, L = x.shape
        nwin = (L - N) // hop + 1
        out = torch.empty((S, nwin, self.model.num_classes), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            for s in range(S):
                _native.check(lib.kws_pcm_to_logits(
                    self._h.ptr, C.c_void_p(x[s].data_ptr()), nwin, N, hop, C.c_void_p(out[s].data_ptr()),
                    _native.current_stream_ptr(self._idx)), ModelError)
        return out[0] if single else out

    def infer_host(self, pcm: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """HOST float32 or int16 ``[B, N]`` (ideally pinned) -> HOST logits; copies pipelined in the library.
        16-bit PCM (the wav sample format) halves the PCIe bytes, which is what bounds this call."""
        lib = _native.lib()
        self._bind()
        if pcm.device.type != "cpu" or pcm.dtype not in (torch.float32, torch.int16) or pcm.dim() != 2:
            raise AudioProcessingError("infer_host expects a float32 (or int16 PCM) [B, N] CPU tensor")
        pcm = pcm.contiguous()
        B, N = pcm.shape
        if out is None:
            out = torch.empty((B, self.model.num_classes), dtype=torch.float32, pin_memory=True)
        with torch.cuda.device(self.device):
            fn = lib.kws_pcm_to_logits_host_i16 if pcm.dtype == torch.int16 else lib.kws_pcm_to_logits_host
            _native.check(fn(self._h.ptr, C.c_void_p(pcm.data_ptr()), B, N, C.c_void_p(out.data_ptr())), ModelError)
        return out
