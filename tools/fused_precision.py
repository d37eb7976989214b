"""Which parameter group dominates the fused kernel's tf32 error?  fused vs float64 oracle, F=40, with bias groups zeroed."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import kws_b200  # noqa
from kws_b200.models import DepthwiseSeparableConv
from oracle import dscnn_ref

F = int(sys.argv[1]) if len(sys.argv) > 1 else 40
x = torch.randn(64, 1, 99, F, generator=torch.Generator().manual_seed(1)) * 5.0
for name, keep in (("all biases", lambda k: True), ("no biases", lambda k: False), ("conv1 bias only", lambda k: k.startswith("conv1")),
                   ("pointwise biases only", lambda k: "pointwise" in k), ("depthwise biases only", lambda k: "depthwise" in k)):
    sd = dscnn_ref.init_state_dict(seed=0)
    for k in sd:
        if k.endswith("bias") and not k.startswith("fc") and not keep(k):
            sd[k] = torch.zeros_like(sd[k])
    m = DepthwiseSeparableConv(); m.load_state_dict(sd); m = m.cuda()
    fused = m(x.cuda()).cpu().numpy()
    m.set_kernel_path("layered")
    layered = m(x.cuda()).cpu().numpy()
    want = dscnn_ref.forward({k: v.double() for k, v in sd.items()}, x.double()).numpy()
    print(f"F={F} {name:24s} fused-oracle {np.abs(fused - want).max():.3e}  layered-oracle {np.abs(layered - want).max():.3e}  max|logit| {np.abs(want).max():.3f}")
