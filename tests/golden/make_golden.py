"""Generate the golden fixtures under tests/golden/ by RUNNING THE REFERENCE.

Run in the build container only (needs /root/reference, which does not exist on
the GPU box):

    python tests/golden/make_golden.py

Imports, by path, the reference's own
  * ``kws/libs/models.py``  (``DepthwiseSeparableConv``, models.py:122-183)
  * ``kws/libs/speech_features/sigproc.py`` (framesig/magspec/powspec/preemphasis)
and stores small input/output vectors.  ``python_speech_features`` is not
installable here, so no psf-generated vector exists (see oracle/psf_mfcc.py).
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


if __name__ == "__main__":
    golden_sigproc()
    golden_model()
    for f in sorted(os.listdir(HERE)):
        print(f, os.path.getsize(os.path.join(HERE, f)))
