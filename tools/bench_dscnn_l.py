"""BASELINE config 2 shape: 40-MFCC + 'DS-CNN-L' (276 channels, 5 blocks; SURVEY a-X1) forward, layered kernels.
usage: bench_dscnn_l.py [B] [reps]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import kws_b200  # noqa: F401
from kws_b200.models import DepthwiseSeparableConv
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
torch.manual_seed(0)
m = DepthwiseSeparableConv(channels=276, num_blocks=5).cuda()
x = torch.randn(B, 1, 99, 40, device="cuda")
for _ in range(2):
    m(x)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    m(x)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
print(json.dumps({"workload": "DS-CNN-L (276 ch, 5 blocks) forward, features [%d,1,99,40]" % B, "ms": ms,
                  "clips_per_s": B / ms * 1e3, "gemm_tflops": B / ms * 1e3 * 907.5e6 / 1e12}))
