"""PCM -> logits as ONE kernel launch (csrc/pcm_logits_fused.cu; north_star "the whole clip stays on-chip"):
every CTA alternates between the MFCC of its next chunk of clips (features into its own L2-resident scratch) and the
network on that chunk.  The logits must equal the two-kernel path (MFCC kernel, then network kernel) bit for bit --
same arithmetic, different plumbing -- for every batch shape and chunk size (one clip per phase up to the whole
range of a CTA), for both feature widths and both sample types, and match the float64 oracle within the north_star
tolerance."""
import os

import numpy as np
import pytest
import torch

from kws_b200.audio_processor import AudioConfig
from kws_b200.models import DepthwiseSeparableConv
from kws_b200.pipeline import KeywordSpotter
from kws_b200.utils import synth
from oracle import dscnn_ref, psf_mfcc

pytestmark = pytest.mark.gpu


def _spotter(numcep, nfilt, seed=0):
    sd = dscnn_ref.init_state_dict(seed=seed)
    m = DepthwiseSeparableConv()
    m.load_state_dict(sd)
    return sd, KeywordSpotter(m.cuda(), AudioConfig(), numcep=numcep, nfilt=nfilt, device="cuda:0")


@pytest.mark.parametrize("numcep,nfilt", [(40, 40), (10, 26)])
@pytest.mark.parametrize("B", [1, 2, 3, 5, 8, 149, 300, 1301])
def test_single_launch_equals_two_kernels(numcep, nfilt, B):
    sd, ks = _spotter(numcep, nfilt, seed=B)
    pcm = synth.parity_pcm(max(B, 8), seed=B)[:B].cuda()
    ks.configure_single_launch(min_clips=0)                  # two kernels back to back
    l0 = ks.model.launch_count()
    two = ks(pcm)
    assert ks.model.launch_count() - l0 == 2
    for chunk in (32, 1, 6, 4096):                           # many phase switches per CTA ... one chunk per CTA
        ks.configure_single_launch(min_clips=1, chunk_clips=chunk)
        l0 = ks.model.launch_count()
        one = ks(pcm)
        assert ks.model.launch_count() - l0 == 1             # ONE kernel
        assert torch.equal(one, two), (chunk, (one - two).abs().max().item())
    assert torch.equal(ks(pcm), two)                         # launch after launch on the same scratch


def test_single_launch_int16():
    sd, ks = _spotter(40, 40, seed=3)
    g = torch.Generator().manual_seed(4)
    i16 = torch.randint(-32768, 32767, (2050, 16000), generator=g, dtype=torch.int16).cuda()
    ks.configure_single_launch(min_clips=0)
    two = ks(i16)
    ks.configure_single_launch(min_clips=1, chunk_clips=32)
    assert torch.equal(ks(i16), two)
    f32 = i16.float() / 32768.0
    assert torch.equal(ks(f32), two)                         # x/32768 is exact: same bits as the float32 call


def test_single_launch_against_oracle_default_threshold():
    """Default configuration (single launch at every batch size); parity set inside a big batch vs the float64 oracle."""
    sd, ks = _spotter(40, 40, seed=5)
    pcm = synth.uniform_pcm(8192, seed=6)
    pcm[:8] = synth.parity_pcm(8)
    l0 = ks.model.launch_count()
    got = ks(pcm.cuda()).cpu().numpy()
    assert ks.model.launch_count() - l0 == 1
    idx = list(range(8)) + [100, 1023, 1024, 8191]
    feats = psf_mfcc.extract_features_batch(pcm[idx].numpy(), numcep=40, nfilt=40)
    want = dscnn_ref.forward({k: v.double() for k, v in sd.items()}, torch.from_numpy(feats).unsqueeze(1)).numpy()
    np.testing.assert_allclose(got[idx], want, rtol=0, atol=1e-3)
    assert (got[idx].argmax(1) == want.argmax(1)).all()


def test_single_launch_sliding_windows():
    """Overlapping rows (sliding windows of one recording, f4) through the single launch."""
    sd, ks = _spotter(40, 40, seed=7)
    rec = synth.uniform_pcm(1, n=16000 + 160 * 1199, seed=8)[0].cuda()
    ks.configure_single_launch(min_clips=0)
    two = ks.sliding_logits(rec, hop_samples=160)
    ks.configure_single_launch(min_clips=1)
    one = ks.sliding_logits(rec, hop_samples=160)
    assert one.shape == (1200, 12) and torch.equal(one, two)


def test_repeatable_bits_large_batch():
    """The network kernel's compute warps run up to one row step apart (they only ARRIVE at the phase barriers) and
    the single launch alternates phases on shared buffers: a race there would show up as run-to-run differences.
    Twelve runs of a 20 000-clip batch, both paths, must give identical bits, equal to the two-kernel result."""
    sd, ks = _spotter(40, 40, seed=9)
    pcm = synth.uniform_pcm(20000, seed=10).cuda()
    ks.configure_single_launch(min_clips=0)
    ref = ks(pcm).clone()
    feats = None
    for rep in range(6):
        assert torch.equal(ks(pcm), ref), ("two kernels", rep)
    ks.configure_single_launch(min_clips=1, chunk_clips=32)
    for rep in range(6):
        assert torch.equal(ks(pcm), ref), ("single launch", rep)
    del feats
