"""Stock-PyTorch eager baseline on the SAME B200 ("the real bar to beat").

Nothing of this repository's kernels is on this path: features are torch ops (unfold / torch.fft.rfft /
matmul), the network is stock ``nn.Conv2d`` modules in the reference's architecture
(kws/libs/models.py:122-183, restated here), training is autograd + ``torch.optim.Adam``.  Printed next
to the repo's own numbers for the same workloads so the speed-up over "just move the reference to the
GPU" is on record.  usage: bench_eager.py [clips] [F]
"""
import json
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.nn as nn  # noqa: E402
import torch.nn.functional as Fn  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
F = int(sys.argv[2]) if len(sys.argv) > 2 else 40
dev = torch.device("cuda", 0)
torch.manual_seed(0)
N, FL, HOP, NFFT = 16000, 400, 160, 512


# ---- eager MFCC (psf 0.6 semantics, fp32): the straightforward torch translation ----
def mel_fb(nfilt):
    hz2mel = lambda hz: 2595.0 * math.log10(1 + hz / 700.0)
    pts = torch.linspace(hz2mel(0.0), hz2mel(8000.0), nfilt + 2, dtype=torch.float64)
    bins = torch.floor((NFFT + 1) * (700.0 * (10 ** (pts / 2595.0) - 1)) / 16000.0).long().tolist()
    fb = torch.zeros(nfilt, NFFT // 2 + 1, dtype=torch.float64)
    for j in range(nfilt):
        for i in range(bins[j], bins[j + 1]):
            fb[j, i] = (i - bins[j]) / (bins[j + 1] - bins[j])
        for i in range(bins[j + 1], bins[j + 2]):
            fb[j, i] = (bins[j + 2] - i) / (bins[j + 2] - bins[j + 1])
    return fb


def dct_mat(nfilt, ncep):
    n = torch.arange(nfilt, dtype=torch.float64)
    k = torch.arange(ncep, dtype=torch.float64)[:, None]
    m = torch.cos(math.pi * k * (2 * n + 1) / (2 * nfilt)) * math.sqrt(2.0 / nfilt)
    m[0] *= 1 / math.sqrt(2.0)
    return m


FB = mel_fb(F).float().to(dev).t().contiguous()
DCT = dct_mat(F, F).float().to(dev).t().contiguous()
LIFT = (1 + 11.0 * torch.sin(math.pi * torch.arange(F, dtype=torch.float64) / 22.0)).float().to(dev)
EPS = 2.220446049250313e-16


def eager_mfcc(pcm):
    y = torch.cat([pcm[:, :1], pcm[:, 1:] - 0.97 * pcm[:, :-1]], dim=1)
    nf = 1 + math.ceil((N - FL) / HOP)
    y = Fn.pad(y, (0, (nf - 1) * HOP + FL - N))
    frames = y.unfold(1, FL, HOP)
    spec = torch.fft.rfft(frames, n=NFFT, dim=-1)
    p = (spec.real ** 2 + spec.imag ** 2) / NFFT
    energy = p.sum(-1)
    energy = torch.where(energy == 0, torch.full_like(energy, EPS), energy)
    mel = p @ FB
    mel = torch.where(mel == 0, torch.full_like(mel, EPS), mel)
    c = (torch.log(mel) @ DCT) * LIFT
    c[..., 0] = torch.log(energy)
    return c


class Block(nn.Module):
    def __init__(self, ch):
        super().__init__()
        self.depthwise = nn.Conv2d(ch, ch, 3, 1, 1, groups=ch)
        self.pointwise = nn.Conv2d(ch, ch, 1, 1, 1)

    def forward(self, x):
        return torch.relu(self.pointwise(self.depthwise(x)))


class EagerDSCNN(nn.Module):
    def __init__(self, ncls=12, ch=64, nb=4):
        super().__init__()
        self.conv1 = nn.Conv2d(1, ch, 10, 2, 2)
        self.blocks = nn.Sequential(*[Block(ch) for _ in range(nb)])
        self.fc = nn.Linear(ch, ncls)

    def forward(self, x):
        x = self.blocks(torch.relu(self.conv1(x)))
        return self.fc(Fn.adaptive_avg_pool2d(x, 1).flatten(1))


def timed(fn, reps=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


g = torch.Generator(device=dev).manual_seed(0)
pcm = torch.rand((B, N), device=dev, generator=g) * 2 - 1
CH = 2048  # clips per eager call (the unfold/rfft temporaries are ~20x the input)
out = {"clips": B, "F": F, "torch": torch.__version__}


def mfcc_all():
    for i in range(0, B, CH):
        eager_mfcc(pcm[i:i + CH])


with torch.no_grad():
    ms = timed(mfcc_all, reps=3)
    out["eager_mfcc_clips_per_s"] = B / (ms * 1e-3)
    feats = torch.cat([eager_mfcc(pcm[i:i + CH]) for i in range(0, B, CH)]).unsqueeze(1).contiguous()
    model = EagerDSCNN().to(dev).eval()
    for tf32 in (False, True):
        torch.backends.cudnn.allow_tf32 = tf32
        torch.backends.cuda.matmul.allow_tf32 = tf32
        torch.backends.cudnn.benchmark = True

        def fwd_all():
            for i in range(0, B, CH):
                model(feats[i:i + CH])
        ms = timed(fwd_all, reps=3)
        out["eager_model_clips_per_s_%s" % ("tf32" if tf32 else "fp32")] = B / (ms * 1e-3)

    def e2l():
        for i in range(0, B, CH):
            model(eager_mfcc(pcm[i:i + CH]).unsqueeze(1))
    ms = timed(e2l, reps=3)
    out["eager_pcm_to_logits_clips_per_s_tf32"] = B / (ms * 1e-3)

# ---- training step: autograd + Adam, the reference loop body (kws/libs/training.py:285-297) ----
for Ft, Bt in ((10, 1028), (40, 1028)):
    tm = EagerDSCNN().to(dev).train()
    opt = torch.optim.Adam(tm.parameters(), lr=1e-3)
    crit = nn.CrossEntropyLoss()
    x = torch.randn(Bt, 1, 99, Ft, device=dev)
    y = torch.randint(0, 12, (Bt,), device=dev)

    def step():
        opt.zero_grad()
        loss = crit(tm(x), y)
        loss.backward()
        opt.step()
    ms = timed(step, reps=10, warm=3)
    out["eager_train_samples_per_s_F%d_B%d_tf32" % (Ft, Bt)] = Bt / (ms * 1e-3)

# ---- this repository, same workloads, same process ----
import kws_b200  # noqa: F401,E402
from kws_b200.audio_processor import AudioConfig, AudioProcessor  # noqa: E402
from kws_b200.models import DepthwiseSeparableConv  # noqa: E402
from kws_b200.pipeline import KeywordSpotter  # noqa: E402

proc = AudioProcessor(None, AudioConfig(), device=dev)
ours = DepthwiseSeparableConv().to(dev)
ours.load_state_dict({k.replace("blocks.0", "dsconv1").replace("blocks.1", "dsconv2").replace("blocks.2", "dsconv3")
                      .replace("blocks.3", "dsconv4"): v for k, v in model.state_dict().items()})
ks = KeywordSpotter(ours, AudioConfig(), numcep=F, nfilt=F, device=dev)
ms = timed(lambda: proc.extract_features(pcm, numcep=F, nfilt=F), reps=5)
out["kws_b200_mfcc_clips_per_s"] = B / (ms * 1e-3)
mine = proc.extract_features(pcm, numcep=F, nfilt=F)
out["mfcc_max_abs_diff_vs_eager"] = float((mine - feats[:, 0]).abs().max())
ms = timed(lambda: ours(feats), reps=5)
out["kws_b200_model_clips_per_s"] = B / (ms * 1e-3)
with torch.no_grad():
    torch.backends.cudnn.allow_tf32 = False
    ref_logits = torch.cat([model(feats[i:i + CH]) for i in range(0, B, CH)])
out["logits_max_abs_diff_vs_eager_fp32"] = float((ours(feats) - ref_logits).abs().max())
ms = timed(lambda: ks(pcm), reps=5)
out["kws_b200_pcm_to_logits_clips_per_s"] = B / (ms * 1e-3)
for k in ("mfcc", "model", "pcm_to_logits"):
    e = out.get("eager_%s_clips_per_s" % k) or out.get("eager_%s_clips_per_s_tf32" % k)
    out["speedup_%s_vs_eager" % k] = out["kws_b200_%s_clips_per_s" % k] / e
print(json.dumps(out))
