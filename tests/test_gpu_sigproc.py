"""GPU parity of the sigproc helpers against the oracle restatement of the reference's torch
module and against the golden vectors produced by the reference itself."""
import os

import numpy as np
import pytest
import torch

from kws_b200.common.errors import AudioProcessingError
from kws_b200.speech_features import sigproc
from oracle import sigproc_ref

pytestmark = pytest.mark.gpu


def _seed0():
    np.random.seed(0)
    return np.random.rand(16000)


def test_reference_test_magspec(golden_dir):
    """The reference's only test (tests/kws/libs/speech_features/test_sigproc.py:8-21), with
    its tolerance, against the golden output of the reference module."""
    g = np.load(os.path.join(golden_dir, "sigproc_seed0.npz"))
    sig = torch.tensor(_seed0()).cuda()
    frames = sigproc.framesig(sig, 400, 160)
    assert frames.dtype == torch.float64 and tuple(frames.shape) == (99, 400)
    np.testing.assert_allclose(frames.cpu().numpy()[g["frame_rows"]], g["frames_hann_rows"], rtol=0, atol=1e-7)
    mag = sigproc.magspec(frames, 512)
    assert tuple(mag.shape) == (99, 257)
    np.testing.assert_allclose(mag.cpu().numpy(), g["magspec_hann"], rtol=1e-5, atol=1e-8)
    pw = sigproc.powspec(frames, 512)
    np.testing.assert_allclose(pw.cpu().numpy()[g["frame_rows"]], g["powspec_hann_rows"], rtol=1e-5, atol=1e-8)
    rect = sigproc.framesig(sig, 400, 160, winfunc=torch.ones)
    np.testing.assert_array_equal(rect.cpu().numpy()[g["frame_rows"]], g["frames_rect_rows"])
    pre = sigproc.preemphasis(sig).cpu().numpy()
    np.testing.assert_array_equal(pre[:64], g["preemph_head"])
    np.testing.assert_array_equal(pre[-64:], g["preemph_tail"])


@pytest.mark.parametrize("n", [1, 399, 400, 401, 560, 561, 16000, 16001])
def test_framesig_shapes_and_indexing(golden_dir, n):
    g = np.load(os.path.join(golden_dir, "framesig_shapes.npz"))
    x = torch.arange(n, dtype=torch.float64).cuda()
    f = sigproc.framesig(x, 400, 160, winfunc=torch.ones)
    assert tuple(f.shape) == tuple(g[str(n)])
    np.testing.assert_array_equal(f.cpu().numpy(), sigproc_ref.framesig(np.arange(n, dtype=np.float64), 400, 160,
                                                                        window=np.ones(400)))


@pytest.mark.parametrize("dtype,rtol,atol", [(torch.float32, 2e-4, 2e-4), (torch.float64, 1e-9, 1e-10)])
@pytest.mark.parametrize("nfft", [512, 256, 1024, 400, 640, 300])
def test_spec_against_oracle(dtype, rtol, atol, nfft):
    rs = np.random.RandomState(1)
    frames = rs.randn(37, 400)
    t = torch.tensor(frames, dtype=dtype).cuda()
    want_m = sigproc_ref.magspec(frames.astype(np.float64 if dtype == torch.float64 else np.float32).astype(np.float64), nfft)
    np.testing.assert_allclose(sigproc.magspec(t, nfft).cpu().numpy(), want_m, rtol=rtol, atol=atol)
    np.testing.assert_allclose(sigproc.powspec(t, nfft).cpu().numpy(), want_m ** 2 / nfft, rtol=rtol * 2, atol=atol)


def test_batched_and_float32_preemphasis():
    g = torch.Generator().manual_seed(2)
    x = torch.rand(5, 1000, generator=g) - 0.5
    y = sigproc.preemphasis(x.cuda(), 0.97).cpu()
    ref = torch.cat([x[:, :1], x[:, 1:] - 0.97 * x[:, :-1]], dim=1)
    assert torch.equal(y, ref)                                  # bit-exact (two roundings, no FMA)
    f = sigproc.framesig(x.cuda(), 400, 160)
    assert f.dtype == torch.float32 and tuple(f.shape) == (5, 5, 400)
    one = sigproc.framesig(x[2].cuda(), 400, 160)
    assert torch.equal(f[2], one)
    want = sigproc_ref.framesig(x[2].numpy(), 400, 160)
    np.testing.assert_allclose(one.cpu().numpy(), want, rtol=0, atol=1e-7)
    m = sigproc.magspec(f, 512)
    assert tuple(m.shape) == (5, 5, 257)


def test_errors():
    with pytest.raises(AudioProcessingError, match="no CPU fallback"):
        sigproc.preemphasis(torch.zeros(10))
    with pytest.raises(AudioProcessingError):
        sigproc.framesig(torch.zeros(10, dtype=torch.int32).cuda(), 4, 2)
