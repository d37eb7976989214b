"""kws_b200 -- B200-native (sm_100a) hot path of z430/keyword-spotting.

PCM -> MFCC -> depthwise-separable KWS network -> logits, behind the reference's Python call
signatures.  All compute is in ``libkws_b200.so`` (hand-written CUDA behind a C ABI, see
``include/kws_b200.h``); this package is the thin host-side mirror of the reference interface.
Import as ``import kws_b200`` (root-level alias module) because the directory name has a hyphen.
"""
from .common.errors import (AudioProcessingError, DatasetError, KWSError, ModelError,  # noqa: F401
                            NativeLibraryError, handle_error)

__version__ = "0.1.0"


def __getattr__(name):
    # lazy: importing the package must not require torch/CUDA to be initialised
    if name in ("AudioConfig", "AudioProcessor"):
        from . import audio_processor
        return getattr(audio_processor, name)
    if name in ("DepthwiseSeparableConv", "DepthwiseSeparableConvBlock", "KeywordSpottingModel"):
        from . import models
        return getattr(models, name)
    if name == "KeywordSpotter":
        from .pipeline import KeywordSpotter
        return KeywordSpotter
    if name == "sigproc":
        from .speech_features import sigproc
        return sigproc
    raise AttributeError(name)
