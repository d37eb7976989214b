"""kws_pcm_to_logits_static: the single-launch PCM -> logits call with its feature scratch in a static __device__
array (no cudaMalloc / cudaFree, no counterpart in the reference).
Called through the C ABI and compared with the existing entry point, kws_pcm_to_logits, on two calls in a row (the
second with other inputs and another batch size: the array keeps the first call's contents).  Collected last on
purpose."""
import ctypes as C

import pytest
import torch

from kws_b200 import _native
from kws_b200.audio_processor import AudioConfig
from kws_b200.models import DepthwiseSeparableConv
from kws_b200.pipeline import KeywordSpotter
from kws_b200.utils import synth
from oracle import dscnn_ref

pytestmark = pytest.mark.gpu
TOL = 1e-3     # north_star logit tolerance, as in tests/test_gpu_model.py; the two paths are in fact bit-identical


def _spotter(numcep, nfilt, seed):
    m = DepthwiseSeparableConv()
    m.load_state_dict(dscnn_ref.init_state_dict(seed=seed))
    ks = KeywordSpotter(m.cuda(), AudioConfig(), numcep=numcep, nfilt=nfilt, device="cuda:0")
    ks._bind()
    return ks


def _call(fn, ks, pcm):
    B, N = pcm.shape
    out = torch.full((B, ks.model.num_classes), float("nan"), dtype=torch.float32, device=pcm.device)
    rc = fn(ks._h.ptr, C.c_void_p(pcm.data_ptr()), B, N, pcm.stride(0), C.c_void_p(out.data_ptr()),
            _native.current_stream_ptr(0))
    torch.cuda.synchronize()
    return rc, out


@pytest.mark.parametrize("numcep,nfilt", [(40, 40), (10, 26)])
def test_static_scratch_entry_point_matches_existing_one(numcep, nfilt):
    lib = _native.lib()
    ks = _spotter(numcep, nfilt, seed=31)
    first = synth.parity_pcm(701, seed=32).cuda()              # odd batch: a half-empty clip group at the end
    second = synth.uniform_pcm(1500, seed=33).cuda() * 0.25     # other inputs, other batch size, same scratch
    for pcm in (first, second):
        rc, want = _call(lib.kws_pcm_to_logits, ks, pcm)
        assert rc == 0, _native.last_error()
        l0 = ks.model.launch_count()
        rc, got = _call(lib.kws_pcm_to_logits_static, ks, pcm)
        assert rc == 0, _native.last_error()
        assert ks.model.launch_count() - l0 == 1                # one kernel, nothing else
        assert torch.isfinite(got).all()
        assert (got - want).abs().max().item() <= TOL
        assert (got.argmax(1) == want.argmax(1)).all()
        assert torch.equal(got, want)                           # same kernel, same arithmetic: same bits
    # a third call repeating the first inputs gives the first result again
    rc, again = _call(lib.kws_pcm_to_logits_static, ks, first)
    rc2, want = _call(lib.kws_pcm_to_logits, ks, first)
    assert rc == 0 and rc2 == 0 and torch.equal(again, want)


def test_static_scratch_rejects_what_does_not_fit():
    lib = _native.lib()
    ks = _spotter(40, 40, seed=34)
    long_clips = torch.zeros(4, 16000 * 3, device="cuda")       # 299 frames x 40: one clip pair per SM exceeds 10 MiB
    rc, _ = _call(lib.kws_pcm_to_logits_static, ks, long_clips)
    assert rc == -1                                             # KWS_ERR_INVALID
    msg = _native.last_error()
    assert "static scratch" in msg and "kws_pcm_to_logits" in msg
    rc, out = _call(lib.kws_pcm_to_logits, ks, long_clips)      # the allocating entry point still takes them
    assert rc == 0 and torch.isfinite(out).all()
    # and the static one keeps working afterwards
    pcm = synth.uniform_pcm(64, seed=35).cuda()
    rc, got = _call(lib.kws_pcm_to_logits_static, ks, pcm)
    rc2, want = _call(lib.kws_pcm_to_logits, ks, pcm)
    assert rc == 0 and rc2 == 0 and torch.equal(got, want)
