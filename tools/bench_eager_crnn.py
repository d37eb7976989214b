"""Stock-PyTorch eager CRNN (Conv2d + cuDNN nn.GRU + Linear, the same definition as kws_b200.models.CRNN) on the
same B200, next to crnn.cu.  usage: bench_eager_crnn.py [clips]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.nn as nn  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
dev = torch.device("cuda", 0)
torch.manual_seed(0)


class EagerCRNN(nn.Module):
    def __init__(self, F=40, C=16, H=64, K=12):
        super().__init__()
        self.conv1 = nn.Conv2d(1, C, 10, 2, 2)
        W1 = (F + 4 - 10) // 2 + 1
        self.gru = nn.GRU(C * W1, H, num_layers=1, batch_first=True)
        self.fc = nn.Linear(H, K)

    def forward(self, x):
        y = torch.relu(self.conv1(x))
        b, c, t, w = y.shape
        seq = y.permute(0, 2, 1, 3).reshape(b, t, c * w)
        out, _ = self.gru(seq)
        return self.fc(out[:, -1])


def timed(fn, reps=5, warm=3):
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


x = torch.randn(B, 1, 99, 40, device=dev)
net = EagerCRNN().to(dev).eval()
out = {"clips": B}
CH = 4096
with torch.no_grad():
    for tf32 in (False, True):
        torch.backends.cudnn.allow_tf32 = tf32
        torch.backends.cuda.matmul.allow_tf32 = tf32

        def run():
            for i in range(0, B, CH):
                net(x[i:i + CH])
        out["eager_crnn_clips_per_s_%s" % ("tf32" if tf32 else "fp32")] = B / (timed(run) * 1e-3)

import kws_b200  # noqa: F401,E402
from kws_b200.models import CRNN  # noqa: E402

ours = CRNN(n_features=40).to(dev)
out["kws_b200_crnn_clips_per_s"] = B / (timed(lambda: ours(x)) * 1e-3)
print(json.dumps(out))
