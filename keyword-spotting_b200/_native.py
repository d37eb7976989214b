"""ctypes binding of ``libkws_b200.so`` (C ABI in ``include/kws_b200.h``).

There is no CPU fallback: if the shared library is missing the import of any compute entry
point raises ``NativeLibraryError`` telling the user to build it
(``python -c "import __graft_entry__ as g; g.build()"`` or
``python keyword-spotting_b200/build.py``).
"""
from __future__ import annotations

import ctypes as C
import os
import threading

from .common.errors import NativeLibraryError

_HERE = os.path.dirname(os.path.abspath(__file__))
# KWS_LIB: an experiment build of the same sources (build.py --out=...), for A/B timing runs only
LIB_PATH = os.path.join(_HERE, os.environ.get("KWS_LIB", "libkws_b200.so"))

_lib = None
_lock = threading.Lock()


class FeatureConfig(C.Structure):
    """Mirror of ``kws_feature_config``."""

    _fields_ = [
        ("sample_rate", C.c_int32),
        ("frame_len", C.c_int32),
        ("frame_step", C.c_int32),
        ("nfft", C.c_int32),
        ("nfilt", C.c_int32),
        ("numcep", C.c_int32),
        ("preemph", C.c_float),
        ("ceplifter", C.c_int32),
        ("append_energy", C.c_int32),
        ("lowfreq", C.c_double),
        ("highfreq", C.c_double),
        ("window_host", C.POINTER(C.c_float)),
    ]


_vp, _i64, _i32, _dbl, _f32 = C.c_void_p, C.c_int64, C.c_int, C.c_double, C.c_float

# name -> (restype, argtypes); every symbol include/kws_b200.h declares
SIGNATURES = {
    "kws_version": (_i32, []),
    "kws_last_error": (C.c_char_p, []),
    "kws_create": (_i32, [_i32, C.POINTER(_vp)]),
    "kws_destroy": (_i32, [_vp]),
    "kws_launch_count": (_i64, [_vp]),
    "kws_probe_fp32_tflops": (_i32, [_vp, C.POINTER(_dbl), C.POINTER(_dbl)]),
    "kws_feature_config_default": (None, [C.POINTER(FeatureConfig)]),
    "kws_round_half_up": (_i64, [_dbl]),
    "kws_num_frames": (_i64, [_i64, _i64, _i64]),
    "kws_mel_bin_edges": (_i32, [_i32, _i32, _i32, _dbl, _dbl, C.POINTER(C.c_int32)]),
    "kws_mel_filterbank": (_i32, [_i32, _i32, _i32, _dbl, _dbl, C.POINTER(_dbl)]),
    "kws_dct_lifter_matrix": (_i32, [_i32, _i32, _i32, C.POINTER(_dbl)]),
    "kws_features_configure": (_i32, [_vp, C.POINTER(FeatureConfig)]),
    "kws_features_ncep": (_i32, [_vp]),
    "kws_mfcc": (_i32, [_vp, _vp, _i64, _i64, _i64, _vp, _vp]),
    "kws_logfbank": (_i32, [_vp, _vp, _i64, _i64, _i64, _vp, _vp]),
    "kws_logfbank_i16": (_i32, [_vp, _vp, _i64, _i64, _i64, _vp, _vp]),
    "kws_mfcc_i16": (_i32, [_vp, _vp, _i64, _i64, _i64, _vp, _vp]),
    "kws_mfcc_host": (_i32, [_vp, _vp, _i64, _i64, _vp]),
    "kws_sigproc_preemphasis": (_i32, [_i32, _vp, _vp, _i64, _i64, _dbl, _vp]),
    "kws_sigproc_framesig": (_i32, [_i32, _vp, _i64, _i64, _i64, _i64, _vp, _vp, _vp]),
    "kws_sigproc_spec": (_i32, [_i32, _vp, _i64, _i64, _i64, _i32, _vp, _vp]),
    "kws_sigproc_launch_count": (_i64, []),
    "kws_model_create": (_i32, [_vp, _i32, _i32, _i32]),
    "kws_model_load_tensor": (_i32, [_vp, C.c_char_p, _vp, C.POINTER(_i64), _i32]),
    "kws_model_finalize": (_i32, [_vp]),
    "kws_model_set_path": (_i32, [_vp, _i32]),
    "kws_mfcc_host_i16": (_i32, [_vp, _vp, _i64, _i64, _vp]),
    "kws_augment": (_i32, [_vp, _vp, _i64, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "kws_stream_edge_frames": (_i32, [_vp, _vp, _i64, _i64, _i64, _i64, _i64, _i64, _f32, _vp, _vp]),
    "kws_stream_gather_windows": (_i32, [_vp, _vp, _i64, _vp, _i64, _i64, _i64, _i64, _vp, _vp]),
    "kws_crnn_create": (_i32, [_vp, _i32, _i32, _i32, _i32, _i32]),
    "kws_crnn_load_tensor": (_i32, [_vp, C.c_char_p, _vp, C.POINTER(_i64), _i32]),
    "kws_crnn_finalize": (_i32, [_vp]),
    "kws_crnn_forward": (_i32, [_vp, _vp, _i64, _i64, _i64, _vp, _vp]),
    "kws_softmax": (_i32, [_vp, _vp, _i64, _i64, _vp, _vp]),
    "kws_train_num_params": (_i64, [_vp]),
    "kws_train_forward_backward": (_i32, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, _vp, _vp, _vp, _vp]),
    "kws_train_adam": (_i32, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _f32, _f32, _f32, _f32, _f32, _vp]),
    "kws_model_forward": (_i32, [_vp, _vp, _i64, _i64, _i64, _vp, _vp]),
    "kws_model_workspace_bytes": (_i64, [_vp, _i64, _i64, _i64]),
    "kws_pcm_to_logits_configure": (_i32, [_vp, _i64, _i32]),
    "kws_pcm_to_logits": (_i32, [_vp, _vp, _i64, _i64, _i64, _vp, _vp]),
    "kws_pcm_to_logits_static": (_i32, [_vp, _vp, _i64, _i64, _i64, _vp, _vp]),
    "kws_pcm_to_logits_host": (_i32, [_vp, _vp, _i64, _i64, _vp]),
    "kws_pcm_to_logits_i16": (_i32, [_vp, _vp, _i64, _i64, _i64, _vp, _vp]),
    "kws_pcm_to_logits_host_i16": (_i32, [_vp, _vp, _i64, _i64, _vp]),
}


def lib() -> C.CDLL:
    """Load the shared library once; raise loudly when it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise NativeLibraryError(
                    f"{LIB_PATH} not found - build it with `python keyword-spotting_b200/build.py` "
                    "(there is no CPU or PyTorch fallback)")
            try:
                handle = C.CDLL(LIB_PATH)
            except OSError as e:  # pragma: no cover
                raise NativeLibraryError(f"cannot load {LIB_PATH}: {e}") from e
            for name, (res, args) in SIGNATURES.items():
                try:
                    fn = getattr(handle, name)
                except AttributeError as e:
                    raise NativeLibraryError(f"{LIB_PATH} does not export {name}") from e
                fn.restype = res
                fn.argtypes = args
            _lib = handle
    return _lib


def last_error() -> str:
    msg = lib().kws_last_error()
    return msg.decode("utf-8", "replace") if msg else ""


def check(rc: int, exc=NativeLibraryError):
    """Turn a negative status into the exception type the reference would raise."""
    if rc < 0:
        raise exc(last_error() or f"status {rc}")
    return rc


class Handle:
    """Owns one ``kws_handle_t`` (one per device)."""

    def __init__(self, device: int = 0):
        self._h = _vp()
        check(lib().kws_create(int(device), C.byref(self._h)))
        self.device = int(device)

    @property
    def ptr(self):
        return self._h

    def launch_count(self) -> int:
        return int(lib().kws_launch_count(self._h))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value and _lib is not None:
            _lib.kws_destroy(self._h)
            self._h = _vp()

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass


_handles = {}


def handle_for(device_index: int) -> Handle:
    """Process-wide handle cache keyed by CUDA device index."""
    h = _handles.get(device_index)
    if h is None:
        h = Handle(device_index)
        _handles[device_index] = h
    return h


def current_stream_ptr(device_index: int):
    import torch

    return _vp(torch.cuda.current_stream(device_index).cuda_stream)
