"""CPU tests of the host side: the C-ABI library loads and exports every symbol the header
declares, and its table builders (mel bank, DCT/lifter, frame rule, rounding) agree with the
oracle.  No GPU compute is called here."""
import ctypes as C
import os
import re

import numpy as np
import pytest
from scipy.fftpack import dct

import kws_b200  # noqa: F401  (root-level alias of keyword-spotting_b200/)
from kws_b200 import _native
from oracle import psf_mfcc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    return _native.lib()


def test_library_exports_every_declared_symbol(lib):
    hdr = open(os.path.join(ROOT, "include", "kws_b200.h")).read()
    declared = set(re.findall(r"\b(kws_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 25
    assert declared == set(_native.SIGNATURES), declared ^ set(_native.SIGNATURES)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.kws_version() == 100


def test_no_cpu_fallback_without_gpu(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    h = C.c_void_p()
    rc = lib.kws_create(0, C.byref(h))
    assert rc == -2 and "no CPU fallback" in _native.last_error()
    from kws_b200.common.errors import NativeLibraryError
    with pytest.raises(NativeLibraryError):
        _native.Handle(0)


def test_frame_rule_and_rounding(lib):
    for n in (1, 399, 400, 401, 560, 561, 15999, 16000, 16001, 32000):
        for fl, st in ((400, 160), (512, 256), (640, 320), (1, 1), (400, 401)):
            assert lib.kws_num_frames(n, fl, st) == psf_mfcc.num_frames(n, fl, st)
    for v in (0.5, 1.5, 2.5, 399.5, 400.49999, 0.025 * 16000, 0.01 * 16000, 0.0305 * 16000, 159.5):
        assert lib.kws_round_half_up(v) == psf_mfcc.round_half_up(v), v


@pytest.mark.parametrize("nfilt", [1, 2, 10, 13, 20, 26, 32, 40, 41, 64, 80])
@pytest.mark.parametrize("sr,low,high", [(16000, 0.0, 0.0), (16000, 300.0, 7600.0), (8000, 0.0, 0.0),
                                         (22050, 20.0, 0.0)])
def test_mel_filterbank_matches_oracle(lib, nfilt, sr, low, high):
    nfft = 512
    edges = (C.c_int32 * (nfilt + 2))()
    assert lib.kws_mel_bin_edges(nfilt, nfft, sr, low, high, edges) == 0
    ref_edges = psf_mfcc.mel_bin_edges(nfilt, nfft, sr, low, high or None)
    assert list(edges) == ref_edges.astype(int).tolist()
    fb = np.zeros((nfilt, nfft // 2 + 1))
    assert lib.kws_mel_filterbank(nfilt, nfft, sr, low, high, fb.ctypes.data_as(C.POINTER(C.c_double))) == 0
    np.testing.assert_array_equal(fb, psf_mfcc.get_filterbanks(nfilt, nfft, sr, low, high or None))


def test_mel_rejects_bad_highfreq(lib):
    edges = (C.c_int32 * 28)()
    assert lib.kws_mel_bin_edges(26, 512, 16000, 0.0, 9000.0, edges) == -1
    assert "highfreq is greater than samplerate/2" in _native.last_error()


@pytest.mark.parametrize("nfilt,ncep,L", [(26, 10, 22), (40, 40, 22), (26, 13, 0), (40, 12, 22), (8, 8, 22)])
def test_dct_lifter_matrix(lib, nfilt, ncep, L):
    m = np.zeros((nfilt, ncep))
    assert lib.kws_dct_lifter_matrix(nfilt, ncep, L, m.ctypes.data_as(C.POINTER(C.c_double))) == 0
    x = np.random.RandomState(0).randn(7, nfilt)
    want = psf_mfcc.lifter(dct(x, type=2, axis=1, norm="ortho")[:, :ncep], L)
    np.testing.assert_allclose(x @ m, want, rtol=1e-12, atol=1e-12)


def test_feature_config_default(lib):
    cfg = _native.FeatureConfig()
    lib.kws_feature_config_default(C.byref(cfg))
    assert (cfg.sample_rate, cfg.frame_len, cfg.frame_step, cfg.nfft, cfg.nfilt, cfg.numcep) == \
        (16000, 400, 160, 512, 26, 10)
    assert abs(cfg.preemph - 0.97) < 1e-7 and cfg.ceplifter == 22 and cfg.append_energy == 1


def test_python_mirror_signatures():
    """Same names / keyword arguments as the reference."""
    import inspect
    from kws_b200.audio_processor import AudioConfig, AudioProcessor
    from kws_b200.models import DepthwiseSeparableConv
    from kws_b200.speech_features import sigproc
    assert list(inspect.signature(AudioProcessor.extract_features).parameters) == \
        ["self", "signal", "samplerate", "numcep", "nfft", "winlen", "winstep", "nfilt"]
    assert list(inspect.signature(AudioProcessor.to_mfcc).parameters) == \
        ["self", "signal", "samplerate", "numcep", "nfft", "winlen", "winstep", "nfilt"]
    assert list(inspect.signature(sigproc.framesig).parameters) == ["signal", "frame_len", "frame_step", "winfunc"]
    assert list(inspect.signature(sigproc.magspec).parameters) == ["frames", "NFFT"]
    assert list(inspect.signature(sigproc.powspec).parameters) == ["frames", "NFFT"]
    assert list(inspect.signature(sigproc.preemphasis).parameters) == ["signal", "coeff"]
    c = AudioConfig()
    assert (c.sample_rate, c.frame_length, c.frame_step, c.num_cepstral_coeffs, c.num_mel_filters,
            c.fft_size, c.desired_samples, c.time_shift) == (16000, 0.025, 0.01, 10, 26, 512, 16000, 1600)
    m = DepthwiseSeparableConv()
    sd = m.state_dict()
    assert sum(v.numel() for v in sd.values()) == 26444
    assert list(sd)[:4] == ["conv1.weight", "conv1.bias", "dsconv1.depthwise.weight", "dsconv1.depthwise.bias"]
    assert tuple(sd["dsconv4.pointwise.weight"].shape) == (64, 64, 1, 1) and tuple(sd["fc.weight"].shape) == (12, 64)
    assert float(sd["conv1.bias"].abs().sum()) == 0.0


def test_state_dict_roundtrip_with_reference_keys(golden_dir, tmp_path):
    import torch
    from kws_b200.models import DepthwiseSeparableConv
    g = np.load(os.path.join(golden_dir, "dscnn_ref.npz"))
    sd = {k[3:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("sd.")}
    m = DepthwiseSeparableConv(num_classes=12)
    m.load_state_dict(sd)                         # reference keys/shapes accepted unchanged
    p = tmp_path / "model.pth"
    m.save(str(p))
    m2 = DepthwiseSeparableConv(num_classes=12)
    m2.load(str(p), device=torch.device("cpu"))
    for k in sd:
        assert torch.equal(m2.state_dict()[k], sd[k])
    from kws_b200.common.errors import ModelError
    with pytest.raises(ModelError, match="Failed to load model from"):
        m2.load(str(tmp_path / "missing.pth"), device=torch.device("cpu"))
    if not torch.cuda.is_available():
        with pytest.raises(ModelError, match="no CPU fallback"):
            m2(torch.zeros(1, 1, 99, 10))


def test_batchnorm_folding_matches_torch_eval():
    """a-X4: conv + BatchNorm2d(eval) == folded conv (what the kernels consume)."""
    import torch
    from kws_b200.utils.bnfold import fold_state_dict
    torch.manual_seed(0)
    conv = torch.nn.Conv2d(8, 8, 3, padding=1, groups=8)
    bn = torch.nn.BatchNorm2d(8)
    with torch.no_grad():
        bn.weight.normal_(1, 0.2); bn.bias.normal_(0, 0.3)
        bn.running_mean.normal_(0, 0.5); bn.running_var.uniform_(0.5, 2.0)
    bn.eval()
    sd = {f"c.{k}": v for k, v in conv.state_dict().items()}
    sd.update({f"b.{k}": v for k, v in bn.state_dict().items()})
    folded = fold_state_dict(sd, {"c": "b"}, eps=bn.eps)
    assert set(folded) == {"c.weight", "c.bias"}
    x = torch.randn(2, 8, 9, 7)
    want = bn(conv(x))
    got = torch.nn.functional.conv2d(x, folded["c.weight"], folded["c.bias"], padding=1, groups=8)
    assert torch.allclose(got, want, atol=1e-5, rtol=1e-5)


def test_bench_reference_arm_prints_one_json_line():
    """bench.py contract: stdout carries exactly ONE JSON line (library chatter is routed to stderr); the
    reference arm is the CPU port on the host cores and needs no GPU."""
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1",
                        "--warmup", "0"], capture_output=True, text=True, timeout=300, cwd=root)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, r.stdout
    line = json.loads(lines[0])
    assert line["impl"] == "reference" and line["unit"] == "clips/s" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"] == {"value": line["value"], "unit": "clips/s", "h2d_bytes_per_step": 0,
                           "d2h_bytes_per_step": 0}
    for key in ("metric", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "dtype",
                "data", "config"):
        assert key in line, key
