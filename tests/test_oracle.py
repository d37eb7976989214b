"""CPU tests: the oracle against the golden vectors produced by the reference itself
(tests/golden/make_golden.py) and against the known-answer facts of SURVEY.md App. A."""
import os

import numpy as np
import torch

from oracle import dscnn_ref, psf_mfcc, sigproc_ref


def _seed0_signal():
    np.random.seed(0)
    return np.random.rand(16000)


def test_sigproc_oracle_matches_reference_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "sigproc_seed0.npz"))
    sig = _seed0_signal()
    rows = g["frame_rows"]
    fr = sigproc_ref.framesig(sig, 400, 160)
    assert tuple(g["frames_shape"]) == fr.shape == (99, 400)
    np.testing.assert_allclose(fr[rows], g["frames_hann_rows"], rtol=0, atol=1e-7)  # 1-ulp window
    fr_rect = sigproc_ref.framesig(sig, 400, 160, window=np.ones(400))
    np.testing.assert_array_equal(fr_rect[rows], g["frames_rect_rows"])   # indexing: bit-exact
    # the reference test's tolerance (test_sigproc.py:21)
    np.testing.assert_allclose(sigproc_ref.magspec(fr, 512), g["magspec_hann"], rtol=1e-5, atol=1e-8)
    np.testing.assert_allclose(sigproc_ref.magspec(fr_rect, 512)[rows], g["magspec_rect_rows"],
                               rtol=1e-5, atol=1e-8)
    np.testing.assert_allclose(sigproc_ref.powspec(fr, 512)[rows], g["powspec_hann_rows"],
                               rtol=1e-5, atol=1e-8)
    pre = sigproc_ref.preemphasis(sig)
    np.testing.assert_allclose(pre[:64], g["preemph_head"], rtol=0, atol=0)
    np.testing.assert_allclose(pre[-64:], g["preemph_tail"], rtol=0, atol=0)


def test_psf_framing_equals_reference_framing(golden_dir):
    """psf framesig (rect window) == reference torch framesig with winfunc=ones; this is the
    intent the reference test pins."""
    g = np.load(os.path.join(golden_dir, "sigproc_seed0.npz"))
    sig = _seed0_signal()
    fr = psf_mfcc.framesig(sig, 400, 160)
    np.testing.assert_array_equal(fr[g["frame_rows"]], g["frames_rect_rows"])
    np.testing.assert_allclose(psf_mfcc.magspec(fr, 512)[g["frame_rows"]], g["magspec_rect_rows"],
                               rtol=1e-5, atol=1e-8)


def test_framesig_shapes_ragged(golden_dir):
    g = np.load(os.path.join(golden_dir, "framesig_shapes.npz"))
    for n, shape in g.items():
        n = int(n)
        x = np.arange(n, dtype=np.float64)
        assert sigproc_ref.framesig(x, 400, 160, window=np.ones(400)).shape == tuple(shape)
        assert psf_mfcc.framesig(x, 400, 160).shape == tuple(shape)


def test_known_answers_appendix_a():
    # frame sizes and count
    assert psf_mfcc.round_half_up(0.025 * 16000) == 400
    assert psf_mfcc.round_half_up(0.01 * 16000) == 160
    assert psf_mfcc.num_frames(16000, 400, 160) == 99
    # mel bin edges
    e26 = "0 2 4 7 10 13 16 20 24 29 34 40 46 53 60 68 77 87 97 109 122 136 152 169 188 209 231 256"
    e40 = ("0 1 2 4 6 8 10 12 14 16 19 21 24 27 30 33 37 41 45 49 54 59 64 69 75 81 88 95 103 110 "
           "119 128 137 148 158 170 182 195 209 224 239 256")
    assert psf_mfcc.mel_bin_edges(26).astype(int).tolist() == [int(v) for v in e26.split()]
    assert psf_mfcc.mel_bin_edges(40).astype(int).tolist() == [int(v) for v in e40.split()]
    for m, nnz in ((26, 459), (40, 454)):
        fb = psf_mfcc.get_filterbanks(m)
        assert fb.shape == (m, 257)
        assert int((fb != 0).sum()) == nnz
        assert fb[:, 0].sum() == 0 and fb[:, 256].sum() == 0
        assert ((fb != 0).sum(axis=0) <= 2).all()
        assert ((fb != 0).sum(axis=1) > 0).all()
    lift = psf_mfcc.lifter_table(10)
    np.testing.assert_allclose(
        lift, [1, 2.5655, 4.0991, 5.5696, 6.9470, 8.2035, 9.3132, 10.2538, 11.0060, 11.5544],
        atol=5e-5)


def test_all_zero_clip_and_shapes():
    z = np.zeros(16000, dtype=np.float32)
    f = psf_mfcc.extract_features(z)
    assert f.shape == (99, 10) and f.dtype == np.float64
    np.testing.assert_allclose(f[:, 0], -36.04365338911715, rtol=0, atol=1e-12)
    assert np.abs(f[:, 1:]).max() < 1e-12   # DCT of a constant row: rounding only
    f40 = psf_mfcc.extract_features(z, numcep=40, nfilt=40)
    assert f40.shape == (99, 40)
    # numcep > nfilt silently yields nfilt columns
    assert psf_mfcc.extract_features(z, numcep=40).shape == (99, 26)
    # nfft argument is ignored (overridden), audio_processor.py:268
    x = (np.random.RandomState(1).rand(16000).astype(np.float32) - 0.5)
    np.testing.assert_array_equal(psf_mfcc.extract_features(x, nfft=1024), psf_mfcc.extract_features(x))


def test_preemphasis_runs_in_float32():
    x = (np.random.RandomState(2).rand(1000).astype(np.float32) - 0.5)
    y = psf_mfcc.preemphasis(x, 0.97)
    assert y.dtype == np.float32
    ref = np.float32(x[1:]) - (np.float32(0.97) * x[:-1]).astype(np.float32)
    np.testing.assert_array_equal(y[1:], ref)
    assert y[0] == x[0]


def test_model_oracle_matches_reference_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "dscnn_ref.npz"))
    sd = {k[3:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("sd.")}
    assert int(g["num_params"]) == 26444 == sum(v.numel() for v in sd.values())
    for tag in ("f10", "f40"):
        x = torch.from_numpy(g[f"{tag}.x"])
        logits, layers = dscnn_ref.forward(sd, x, return_layers=True)
        np.testing.assert_allclose(logits.numpy(), g[f"{tag}.logits"], rtol=1e-5, atol=1e-6)
        assert [list(l.shape[1:]) for l in layers[:5]] == g[f"{tag}.block_shapes"].tolist()
        np.testing.assert_allclose(layers[4].mean(dim=(2, 3)).numpy(), g[f"{tag}.block4_mean"],
                                   rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(layers[1][0, 0].numpy(), g[f"{tag}.block1_clip0_ch0"],
                                   rtol=1e-5, atol=1e-6)
        hw = dscnn_ref.output_hw(99, x.shape[-1])
        assert tuple(layers[4].shape[2:]) == hw
    assert dscnn_ref.output_hw(99, 10) == (55, 11) and dscnn_ref.output_hw(99, 40) == (55, 26)


def test_pointwise_border_is_relu_bias(golden_dir):
    g = np.load(os.path.join(golden_dir, "dscnn_ref.npz"))
    sd = {k[3:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("sd.")}
    x = torch.from_numpy(g["f10.x"])
    _, layers = dscnn_ref.forward(sd, x, return_layers=True)
    b = torch.relu(sd["dsconv1.pointwise.bias"])
    np.testing.assert_allclose(layers[1][0, :, 0, 0].numpy(), b.numpy(), rtol=0, atol=0)
    np.testing.assert_allclose(layers[1][0, :, -1, 2].numpy(), b.numpy(), rtol=0, atol=0)


def test_crnn_oracle_matches_torch_gru():
    """The explicit-matmul GRU of oracle/crnn_ref.py is torch.nn.GRU (the a-X3 definition)."""
    import torch
    from oracle import crnn_ref
    torch.manual_seed(0)
    conv = torch.nn.Conv2d(1, 8, 10, stride=2, padding=2).double()
    gru = torch.nn.GRU(8 * 18, 24, 2, batch_first=True).double()
    fc = torch.nn.Linear(24, 12).double()
    sd = {f"conv1.{k}": v for k, v in conv.state_dict().items()}
    sd.update({f"gru.{k}": v for k, v in gru.state_dict().items()})
    sd.update({f"fc.{k}": v for k, v in fc.state_dict().items()})
    x = torch.randn(3, 1, 99, 40, dtype=torch.float64)
    y = torch.relu(conv(x))
    seq = y.permute(0, 2, 1, 3).reshape(3, y.shape[2], -1)
    want = fc(gru(seq)[0][:, -1])
    got = crnn_ref.forward(sd, x)
    assert torch.allclose(got, want, atol=1e-10)


def test_psf_front_stages_pinned_by_reference_sigproc(golden_dir):
    """psf stages 1-5 (float32 pre-emphasis, float64 framing with a rectangular window, |rFFT|^2/512,
    frame energy) of the restatement against vectors produced by RUNNING the reference's own torch
    sigproc helpers on the parity set (tests/golden/make_golden.py::golden_psf_front)."""
    from kws_b200.utils import synth  # noqa: F401  (pure numpy/torch-CPU generator)
    g = np.load(os.path.join(golden_dir, "psf_front_parity8.npz"))
    pcm = synth.parity_pcm(8).numpy()
    rows = g["frame_rows"]
    for b, x in enumerate(pcm):
        y = psf_mfcc.preemphasis(x, 0.97)
        assert y.dtype == np.float32
        np.testing.assert_allclose(y.astype(np.float64).sum(), g["preemph_sum"][b], rtol=1e-12, atol=1e-12)
        ps = psf_mfcc.powspec(psf_mfcc.framesig(y, 400, 160), 512)
        scale = max(float(g["energy"][b].max()), 1e-30)
        np.testing.assert_allclose(ps[rows], g["pspec_rows"][b], rtol=1e-9, atol=1e-12 * scale)
        np.testing.assert_allclose(ps.sum(axis=1), g["energy"][b], rtol=1e-10, atol=1e-12 * scale)
        # ... and through the restated fbank(): its energy output is that row sum with 0 -> eps
        _, energy = psf_mfcc.fbank(x, nfilt=40)
        want = np.where(g["energy"][b] == 0, psf_mfcc.EPS, g["energy"][b])
        np.testing.assert_allclose(energy, want, rtol=1e-10, atol=1e-12 * scale)


def test_train_step_oracle_is_torch_adam():
    """oracle/dscnn_ref.cross_entropy_train_step re-implements Adam by hand; three of its steps must
    equal three steps of torch.optim.Adam(lr=1e-3) + CrossEntropyLoss, which is what the reference builds
    (kws/libs/training.py:112-123)."""
    sd = dscnn_ref.init_state_dict(seed=4, dtype=torch.float64)
    g = torch.Generator().manual_seed(5)
    xs = torch.randn(3, 6, 1, 99, 10, generator=g, dtype=torch.float64)
    ys = torch.randint(0, 12, (3, 6), generator=g)
    params = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    opt = torch.optim.Adam(list(params.values()), lr=1e-3)
    crit = torch.nn.CrossEntropyLoss()
    cur, state = sd, None
    for i in range(3):
        opt.zero_grad()
        loss_t = crit(dscnn_ref.forward(params, xs[i]), ys[i])
        loss_t.backward()
        opt.step()
        loss_o, _, cur, state = dscnn_ref.cross_entropy_train_step(cur, xs[i], ys[i], state)
        assert abs(float(loss_t) - float(loss_o)) < 1e-12
    for k in sd:
        np.testing.assert_allclose(cur[k].numpy(), params[k].detach().numpy(), rtol=1e-9, atol=1e-12)


def test_train_step_oracle_matches_reference_trainer(golden_dir):
    """The same oracle against one epoch (four batches) of the REFERENCE's own KWSTrainer.train_epoch
    (kws/libs/training.py:266-345), run on CPU by tests/golden/make_golden.py::golden_trainer."""
    g = np.load(os.path.join(golden_dir, "ref_trainer_epoch.npz"))
    sd = {k[4:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("sd0.")}
    want = {k[4:]: g[k] for k in g.files if k.startswith("sd1.")}
    xs, ys = torch.from_numpy(g["x"]), torch.from_numpy(g["y"])
    cur, state, losses = sd, None, []
    for i in range(xs.shape[0]):
        loss, _, cur, state = dscnn_ref.cross_entropy_train_step(cur, xs[i], ys[i], state)
        losses.append(float(loss))
    assert abs(np.mean(losses) - float(g["epoch_loss"])) < 1e-5
    for k in want:   # fp32 CPU training: thread-order rounding only
        np.testing.assert_allclose(cur[k].numpy(), want[k], rtol=2e-4, atol=2e-6)


def test_reference_checkpoint_golden_layout(golden_dir):
    """The checkpoint written by the reference's Trainer.save_checkpoint (training.py:199-216) has the
    layout load_reference_checkpoint() expects and holds the post-epoch parameters."""
    ck = torch.load(os.path.join(golden_dir, "ref_trainer_checkpoint.pth"), map_location="cpu", weights_only=False)
    assert set(ck) == {"model_state_dict", "optimizer_state_dict", "metrics", "config", "epoch"}
    g = np.load(os.path.join(golden_dir, "ref_trainer_epoch.npz"))
    for k, v in ck["model_state_dict"].items():
        np.testing.assert_array_equal(v.numpy(), g["sd1." + k])
    assert ck["optimizer_state_dict"]["param_groups"][0]["lr"] == 1e-3
