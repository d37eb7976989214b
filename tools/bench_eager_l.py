"""Stock-PyTorch eager forward of DS-CNN-L (276 channels, 5 blocks) on the same B200, next to model_gemm.cu.
usage: bench_eager_l.py [clips] [channels] [blocks]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.nn as nn  # noqa: E402
import torch.nn.functional as Fn  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
CH = int(sys.argv[2]) if len(sys.argv) > 2 else 276
NB = int(sys.argv[3]) if len(sys.argv) > 3 else 5
dev = torch.device("cuda", 0)
torch.manual_seed(0)


class Block(nn.Module):
    def __init__(self, ch):
        super().__init__()
        self.depthwise = nn.Conv2d(ch, ch, 3, 1, 1, groups=ch)
        self.pointwise = nn.Conv2d(ch, ch, 1, 1, 1)

    def forward(self, x):
        return torch.relu(self.pointwise(self.depthwise(x)))


class Net(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv1 = nn.Conv2d(1, CH, 10, 2, 2)
        self.blocks = nn.Sequential(*[Block(CH) for _ in range(NB)])
        self.fc = nn.Linear(CH, 12)

    def forward(self, x):
        x = self.blocks(torch.relu(self.conv1(x)))
        return self.fc(Fn.adaptive_avg_pool2d(x, 1).flatten(1))


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
out = {"clips": B, "channels": CH, "blocks": NB}
net = Net().to(dev).eval()
torch.backends.cudnn.benchmark = True
with torch.no_grad():
    for name, tf32, cl in (("fp32", False, False), ("tf32", True, False), ("tf32_channels_last", True, True)):
        torch.backends.cudnn.allow_tf32 = tf32
        torch.backends.cuda.matmul.allow_tf32 = tf32
        m = net.to(memory_format=torch.channels_last) if cl else net
        xi = x.to(memory_format=torch.channels_last) if cl else x
        ms = timed(lambda: m(xi))
        out["eager_%s_clips_per_s" % name] = B / (ms * 1e-3)
    mb = net.to(torch.bfloat16).to(memory_format=torch.channels_last)
    xb = x.to(torch.bfloat16).to(memory_format=torch.channels_last)
    ms = timed(lambda: mb(xb))
    out["eager_bf16_channels_last_clips_per_s"] = B / (ms * 1e-3)

import kws_b200  # noqa: F401,E402
from kws_b200.models import DepthwiseSeparableConv  # noqa: E402

ours = DepthwiseSeparableConv(channels=CH, num_blocks=NB).to(dev)
ms = timed(lambda: ours(x))
out["kws_b200_clips_per_s"] = B / (ms * 1e-3)
print(json.dumps(out))
