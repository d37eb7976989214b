"""Fold an inference-mode BatchNorm2d into the preceding convolution (SURVEY.md a-X4).

The reference network has no BatchNorm (every conv carries a bias, ``kws/libs/models.py:96-106,135``);
BASELINE.json's "BN folded into conv weights" therefore reduces to a host-side weight preparation step:
a conv followed by ``BatchNorm2d`` in eval mode IS a conv with rescaled weights and a new bias, which is
exactly what the kernels consume.  Pure tensor arithmetic on the host, float64 internally."""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch


def fold_batchnorm(conv_weight: torch.Tensor, conv_bias: Optional[torch.Tensor], bn_weight: Optional[torch.Tensor],
                   bn_bias: Optional[torch.Tensor], running_mean: torch.Tensor, running_var: torch.Tensor,
                   eps: float = 1e-5) -> Tuple[torch.Tensor, torch.Tensor]:
    """``BN(conv(x))`` with running statistics == ``conv'(x)``; returns ``(weight', bias')`` in the
    conv's dtype.  ``conv_weight`` is ``[out, in/groups, kh, kw]`` (depthwise and pointwise alike)."""
    w = conv_weight.detach().double()
    out_ch = w.shape[0]
    b = conv_bias.detach().double() if conv_bias is not None else torch.zeros(out_ch, dtype=torch.float64)
    gamma = bn_weight.detach().double() if bn_weight is not None else torch.ones(out_ch, dtype=torch.float64)
    beta = bn_bias.detach().double() if bn_bias is not None else torch.zeros(out_ch, dtype=torch.float64)
    scale = gamma / torch.sqrt(running_var.detach().double() + eps)
    w2 = w * scale.view(-1, *([1] * (w.dim() - 1)))
    b2 = (b - running_mean.detach().double()) * scale + beta
    return w2.to(conv_weight.dtype), b2.to(conv_weight.dtype)


def fold_state_dict(sd: Dict[str, torch.Tensor], pairs: Dict[str, str], eps: float = 1e-5) -> Dict[str, torch.Tensor]:
    """``pairs`` maps a conv prefix to its BatchNorm prefix (e.g. ``{"conv1": "bn1"}``).  Returns a
    state_dict with the reference's key set (conv ``weight``/``bias`` only, BN entries removed)."""
    out = {k: v for k, v in sd.items() if not any(k.startswith(bn + ".") for bn in pairs.values())}
    for conv, bn in pairs.items():
        w, b = fold_batchnorm(sd[f"{conv}.weight"], sd.get(f"{conv}.bias"), sd.get(f"{bn}.weight"), sd.get(f"{bn}.bias"),
                              sd[f"{bn}.running_mean"], sd[f"{bn}.running_var"], eps)
        out[f"{conv}.weight"], out[f"{conv}.bias"] = w, b
    return out
