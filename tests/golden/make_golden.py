"""Generate the golden fixtures under tests/golden/ by RUNNING THE REFERENCE.

Run in the build container only (needs /root/reference, which does not exist on
the GPU box):

    python tests/golden/make_golden.py

Imports, by path, the reference's own
  * ``kws/libs/models.py``  (``DepthwiseSeparableConv``, models.py:122-183)
  * ``kws/libs/speech_features/sigproc.py`` (framesig/magspec/powspec/preemphasis)
  * ``kws/libs/training.py`` (``KWSTrainer.train_epoch`` / ``save_checkpoint``, training.py:199-216,266-345)
and stores small input/output vectors.  ``python_speech_features`` is not
installable here, so no psf-generated vector exists (see oracle/psf_mfcc.py); the psf stages 1-5
(pre-emphasis, framing, |rFFT|^2/N, frame energy) are pinned instead by the reference's own torch
``sigproc`` helpers run on the whole parity set with ``winfunc=torch.ones`` (``golden_psf_front``).
"""
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REF)

from kws.libs.models import DepthwiseSeparableConv  # noqa: E402
from kws.libs.speech_features import sigproc  # noqa: E402

sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))


def golden_sigproc():
    # the reference test's input: tests/kws/libs/speech_features/test_sigproc.py:9-12
    np.random.seed(0)
    signal = np.random.rand(16000)
    st = torch.tensor(signal)
    frames = sigproc.framesig(st, 400, 160)                      # periodic Hann (fp32 values)
    frames_rect = sigproc.framesig(st, 400, 160, winfunc=torch.ones)
    mag = sigproc.magspec(frames, 512)
    mag_rect = sigproc.magspec(frames_rect, 512)
    pw = sigproc.powspec(frames, 512)
    pre = sigproc.preemphasis(st)
    rows = np.array([0, 1, 49, 97, 98])
    np.savez_compressed(
        os.path.join(HERE, "sigproc_seed0.npz"),
        frame_rows=rows,
        frames_hann_rows=frames.numpy()[rows],
        frames_rect_rows=frames_rect.numpy()[rows],
        frames_shape=np.array(frames.shape),
        magspec_hann=mag.numpy(),
        magspec_rect_rows=mag_rect.numpy()[rows],
        powspec_hann_rows=pw.numpy()[rows],
        preemph_head=pre.numpy()[:64],
        preemph_tail=pre.numpy()[-64:],
        preemph_sum=np.array(pre.numpy().sum()),
    )
    # short-signal / ragged edge cases of framesig (frame count rule, sigproc.py:31-40)
    cases = {}
    for n in (1, 399, 400, 401, 560, 561, 16000, 16001):
        f = sigproc.framesig(torch.arange(n, dtype=torch.float64), 400, 160, winfunc=torch.ones)
        cases[str(n)] = np.array(f.shape)
    np.savez(os.path.join(HERE, "framesig_shapes.npz"), **cases)


def golden_model():
    out = {}
    torch.manual_seed(0)
    model = DepthwiseSeparableConv(num_classes=12).eval()
    g = torch.Generator().manual_seed(1)
    with torch.no_grad():
        for name, p in model.named_parameters():
            if name.endswith("bias"):
                p.copy_(torch.randn(p.shape, generator=g) * 0.1)
    for k, v in model.state_dict().items():
        out["sd." + k] = v.numpy()
    for tag, fdim in (("f10", 10), ("f40", 40)):
        x = torch.randn(4, 1, 99, fdim, generator=g)
        acts = []
        hooks = [m.register_forward_hook(lambda _m, _i, o: acts.append(o.detach()))
                 for m in (model.conv1, model.dsconv1, model.dsconv2, model.dsconv3, model.dsconv4)]
        with torch.no_grad():
            logits = model(x)
        for h in hooks:
            h.remove()
        out[f"{tag}.x"] = x.numpy()
        out[f"{tag}.logits"] = logits.numpy()
        # per-layer checks: conv1 is pre-ReLU in the hook (F.relu is applied outside the module)
        out[f"{tag}.conv1_prerelu_clip0"] = acts[0][0].numpy()
        out[f"{tag}.block_shapes"] = np.array([list(a.shape[1:]) for a in acts])
        out[f"{tag}.block4_mean"] = acts[4].mean(dim=(2, 3)).numpy()
        out[f"{tag}.block1_clip0_ch0"] = acts[1][0, 0].numpy()
    out["num_params"] = np.array(sum(p.numel() for p in model.parameters()))
    np.savez_compressed(os.path.join(HERE, "dscnn_ref.npz"), **out)


def golden_psf_front():
    """psf stages 1-5 on the parity set, computed by the REFERENCE's torch sigproc
    (kws/libs/speech_features/sigproc.py:14-103): float32 pre-emphasis (two roundings, like psf's
    numpy step on librosa's float32 audio), then float64 framing with a rectangular window, the
    512-point power spectrum and its row sums (psf's frame energy).  Stored: the full energy
    [8, 99], and the power spectrum of five frames per clip."""
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "kws_synth", os.path.join(os.path.dirname(os.path.dirname(HERE)), "keyword-spotting_b200", "utils", "synth.py"))
    synth = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(synth)
    pcm = synth.parity_pcm(8)
    rows = np.array([0, 1, 49, 97, 98])
    energy, pspec_rows, pre_sum = [], [], []
    for x in pcm:
        y = sigproc.preemphasis(x)                      # float32 in -> float32 out
        assert y.dtype == torch.float32
        fr = sigproc.framesig(y.double(), 400, 160, winfunc=torch.ones)
        ps = sigproc.powspec(fr, 512)
        energy.append(ps.sum(dim=1).numpy())
        pspec_rows.append(ps.numpy()[rows])
        pre_sum.append(float(y.double().sum()))
    np.savez_compressed(os.path.join(HERE, "psf_front_parity8.npz"), frame_rows=rows,
                        energy=np.stack(energy), pspec_rows=np.stack(pspec_rows),
                        preemph_sum=np.array(pre_sum))


def golden_trainer():
    """Run the REFERENCE trainer (kws/libs/training.py: KWSTrainer.train_epoch :266-345 with
    CrossEntropyLoss + torch.optim.Adam built in _setup_training :112-123) for one epoch of four
    batches on synthetic data, CPU, and keep the inputs, the per-epoch loss, the parameters before
    and after, and a checkpoint written by the reference's own Trainer.save_checkpoint (:199-216)."""
    import tempfile

    from kws.libs.training import KWSTrainer, TrainingConfig

    torch.manual_seed(0)
    model = DepthwiseSeparableConv(num_classes=12)
    g = torch.Generator().manual_seed(3)
    with torch.no_grad():
        for name, p in model.named_parameters():
            if name.endswith("bias"):
                p.copy_(torch.randn(p.shape, generator=g) * 0.1)
    sd0 = {k: v.clone() for k, v in model.state_dict().items()}
    nb, bs = 4, 16
    xs = torch.randn(nb, bs, 1, 99, 10, generator=g)
    ys = torch.randint(0, 12, (nb, bs), generator=g)
    loader = [(xs[i], ys[i]) for i in range(nb)]
    with tempfile.TemporaryDirectory() as tmp:
        cfg = TrainingConfig(epochs=1, learning_rate=1e-3, batch_size=bs, checkpoint_dir=tmp)
        torch.set_num_threads(1)
        tr = KWSTrainer(model, loader, loader, cfg)
        assert tr.device.type == "cpu"
        loss, acc = tr.train_epoch(0)
        path = tr.save_checkpoint("checkpoint_epoch_1.pth")
        import shutil
        shutil.copy(path, os.path.join(HERE, "ref_trainer_checkpoint.pth"))
    out = {"x": xs.numpy(), "y": ys.numpy(), "epoch_loss": np.array(loss), "epoch_acc": np.array(acc)}
    for k, v in sd0.items():
        out["sd0." + k] = v.numpy()
    for k, v in model.state_dict().items():
        out["sd1." + k] = v.detach().numpy()
    np.savez_compressed(os.path.join(HERE, "ref_trainer_epoch.npz"), **out)


if __name__ == "__main__":
    golden_sigproc()
    golden_model()
    golden_psf_front()
    golden_trainer()
    for f in sorted(os.listdir(HERE)):
        print(f, os.path.getsize(os.path.join(HERE, f)))
