"""GPU: sliding-window inference over long recordings (the batch form of the reference's
1 s window concept, kws/inference/inference_local.py:114-188).  The windows are read in place as
overlapping rows, so every row must equal the per-window call bit for bit, and a few windows are
checked against the oracles (psf-0.6 MFCC restatement, float64 network)."""
import numpy as np
import pytest
import torch

from kws_b200.audio_processor import AudioConfig, AudioProcessor
from kws_b200.common.errors import AudioProcessingError
from kws_b200.models import DepthwiseSeparableConv
from kws_b200.pipeline import KeywordSpotter
from oracle import dscnn_ref, psf_mfcc

pytestmark = pytest.mark.gpu

N = 16000


def _recording(L, seed):
    g = torch.Generator().manual_seed(seed)
    t = torch.arange(L, dtype=torch.float64) / 16000.0
    x = 0.3 * torch.sin(2 * np.pi * 440.0 * t) * (1.0 + 0.5 * torch.sin(2 * np.pi * 1.5 * t))
    return (x + 0.05 * torch.randn(L, generator=g, dtype=torch.float64)).float()


@pytest.mark.parametrize("hop", [160, 1600, 250, 16000])   # 250: not a multiple of 4 -> the generic kernel
def test_sliding_features_equal_per_window_calls(hop):
    proc = AudioProcessor(None, AudioConfig(), device="cuda:0")
    rec = _recording(N + 37 * 160 + 123, seed=hop).cuda()
    got = proc.sliding_features(rec, hop, numcep=40, nfilt=40)
    nwin = (rec.numel() - N) // hop + 1
    assert tuple(got.shape) == (nwin, 99, 40)
    wins = torch.stack([rec[k * hop:k * hop + N] for k in range(nwin)])
    want = proc.extract_features(wins, numcep=40, nfilt=40)
    if hop % 4 == 0:   # same (warp-specialised) kernel on both sides: identical bits
        assert torch.equal(got, want)
    else:              # unaligned rows run on the generic kernel: equal within the parity tolerance
        assert ((got - want).abs() <= 1e-4 + 1e-3 * want.abs()).all()
    # two windows against the float64 oracle (north_star tolerance)
    for k in (0, nwin - 1):
        ref = psf_mfcc.extract_features_batch(wins[k:k + 1].cpu().numpy(), numcep=40, nfilt=40)[0]
        err = np.abs(got[k].cpu().numpy().astype(np.float64) - ref)
        assert (err <= 1e-4 + 1e-3 * np.abs(ref)).all()


@pytest.mark.parametrize("hop,S", [(1600, 1), (800, 3)])
def test_sliding_logits_equal_per_window_calls(hop, S):
    sd = dscnn_ref.init_state_dict(seed=11)
    m = DepthwiseSeparableConv()
    m.load_state_dict(sd)
    ks = KeywordSpotter(m.cuda(), AudioConfig(), numcep=40, nfilt=40, device="cuda:0")
    L = N + 20 * hop + 400
    rec = torch.stack([_recording(L, seed=100 + s) for s in range(S)]).cuda()
    got = ks.sliding_logits(rec if S > 1 else rec[0], hop)
    nwin = (L - N) // hop + 1
    assert tuple(got.shape) == ((S, nwin, 12) if S > 1 else (nwin, 12))
    got = got.reshape(S, nwin, 12)
    for s in range(S):
        wins = torch.stack([rec[s, k * hop:k * hop + N] for k in range(nwin)])
        assert torch.equal(got[s], ks(wins))
    # one window end to end against the oracles: psf MFCC (float64) -> float64 network
    feats = psf_mfcc.extract_features_batch(rec[0, hop:hop + N].cpu().numpy()[None], numcep=40, nfilt=40)
    want = dscnn_ref.forward({k: v.double() for k, v in sd.items()}, torch.from_numpy(feats)[:, None]).numpy()[0]
    np.testing.assert_allclose(got[0, 1].cpu().numpy(), want, rtol=1e-3, atol=1e-3)
    assert got[0, 1].argmax().item() == want.argmax()


def test_sliding_rejects_short_recordings():
    proc = AudioProcessor(None, AudioConfig(), device="cuda:0")
    with pytest.raises(AudioProcessingError):
        proc.sliding_features(torch.zeros(N - 1).cuda(), 160)
    with pytest.raises(AudioProcessingError):
        proc.sliding_features(torch.zeros(2 * N).cuda(), 0)


@pytest.mark.parametrize("hop,extra", [(160, 123), (320, 0), (1600, 77), (16000, 5)])
def test_incremental_features_share_frames(hop, extra):
    """f4: frames shared by consecutive windows are computed once (the recording as one long clip + two edge frames
    per window + a gather).  Interior frames equal the per-window result bit for bit (same kernel, same frame
    samples); the two edge frames come from a second plan (frame step = frame length, pre-emphasised by
    kws_stream_edge_frames) and hold the parity tolerance; the whole map holds it against the float64 oracle."""
    proc = AudioProcessor(None, AudioConfig(), device="cuda:0")
    rec = _recording(N + 23 * hop + extra, seed=hop + 1).cuda()
    full = proc.sliding_features(rec, hop, numcep=40, nfilt=40)
    l0 = proc.launch_count()
    inc = proc.sliding_features_incremental(rec, hop, numcep=40, nfilt=40)
    assert proc.launch_count() - l0 == 4           # recording MFCC, edge frames, edge MFCC, gather
    assert inc.shape == full.shape == (24, 99, 40)
    assert torch.equal(inc[:, 1:98], full[:, 1:98])
    edge_err = (inc - full).abs()
    assert (edge_err <= 1e-4 + 1e-3 * full.abs()).all(), edge_err.max().item()
    for k in (0, 11, 23):
        ref = psf_mfcc.extract_features_batch(rec[k * hop:k * hop + N].cpu().numpy()[None], numcep=40, nfilt=40)[0]
        err = np.abs(inc[k].cpu().numpy().astype(np.float64) - ref)
        assert (err <= 1e-4 + 1e-3 * np.abs(ref)).all(), (k, err.max())


def test_incremental_logits_and_default_shape():
    sd = dscnn_ref.init_state_dict(seed=13)
    m = DepthwiseSeparableConv()
    m.load_state_dict(sd)
    ks = KeywordSpotter(m.cuda(), AudioConfig(), device="cuda:0")     # the reference's default 10 x 26 features
    rec = _recording(N + 40 * 480, seed=3).cuda()
    want = ks.sliding_logits(rec, 480)
    got = ks.sliding_logits_incremental(rec, 480)
    assert got.shape == want.shape == (41, 12)
    np.testing.assert_allclose(got.cpu().numpy(), want.cpu().numpy(), rtol=0, atol=1e-4)
    assert (got.argmax(1) == want.argmax(1)).all()
    proc = AudioProcessor(None, AudioConfig(), device="cuda:0")
    with pytest.raises(AudioProcessingError):
        proc.sliding_features_incremental(rec, 250)       # not a multiple of the frame step
