"""GPU parity of the CRNN / GRU extension (SURVEY a-X3, BASELINE config 4) against its PyTorch oracle.
Parity is unpinned by the reference (it has no recurrent model); tolerance = north_star's 1e-3 on logits."""
import numpy as np
import pytest
import torch

from kws_b200.audio_processor import AudioConfig, AudioProcessor
from kws_b200.common.errors import ModelError
from kws_b200.models import CRNN
from kws_b200.utils import synth
from oracle import crnn_ref, psf_mfcc

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("F,Cc,H,L,B,T", [(40, 16, 64, 1, 19, 99), (40, 8, 32, 2, 5, 99), (10, 16, 64, 2, 33, 49),
                                          (26, 4, 128, 1, 3, 31), (40, 16, 64, 2, 150, 31), (40, 16, 64, 1, 257, 20),
                                          # tensor-core path edges: one output step, exactly one / one-plus-one clip groups
                                          (40, 16, 64, 1, 1, 6), (40, 16, 64, 1, 128, 11), (40, 16, 64, 2, 129, 13)])
def test_crnn_matches_oracle(F, Cc, H, L, B, T):
    torch.manual_seed(F + Cc + H + L)
    m = CRNN(num_classes=12, n_features=F, conv_channels=Cc, hidden=H, num_layers=L)
    with torch.no_grad():
        m.conv1.bias.normal_(0, 0.1)
        m.fc.weight.normal_(0, 0.2)
        m.fc.bias.normal_(0, 0.1)
    sd = {k: v.detach().double() for k, v in m.state_dict().items()}
    x = torch.randn(B, 1, T, F, generator=torch.Generator().manual_seed(B)) * 3.0
    want = crnn_ref.forward(sd, x.double()).numpy()
    got = m.cuda()(x.cuda()).cpu().numpy()
    np.testing.assert_allclose(got, want, rtol=1e-3, atol=1e-4)
    assert (got.argmax(1) == want.argmax(1)).all()
    assert m.launch_count() > 0


def test_logmel_to_crnn_pipeline():
    """Config 4 shape: PCM -> log-mel (kws_logfbank) -> CRNN -> logits."""
    torch.manual_seed(1)
    m = CRNN(n_features=40)
    with torch.no_grad():
        m.fc.weight.normal_(0, 0.2)
    sd = {k: v.detach().double() for k, v in m.state_dict().items()}
    pcm = synth.parity_pcm(16)
    proc = AudioProcessor(None, AudioConfig(), device="cuda:0")
    feats = proc.extract_logfbank(pcm.cuda(), nfilt=40)
    got = m.cuda()(feats.unsqueeze(1)).cpu().numpy()
    ref_feats = np.stack([psf_mfcc.logfbank(x, nfilt=40) for x in pcm.numpy()])
    want = crnn_ref.forward(sd, torch.from_numpy(ref_feats).unsqueeze(1)).numpy()
    np.testing.assert_allclose(got, want, rtol=1e-3, atol=1e-3)


def test_crnn_errors():
    m = CRNN(n_features=40).cuda()
    with pytest.raises(ModelError, match="conv columns"):
        m(torch.zeros(1, 1, 99, 10).cuda())
    with pytest.raises(ModelError, match="no CPU fallback"):
        m(torch.zeros(1, 1, 99, 40))
    assert m(torch.zeros(0, 1, 99, 40).cuda()).shape == (0, 12)
