"""Model oracle: plain PyTorch (CPU) restatement of the reference KWS network.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

Restates ``DepthwiseSeparableConv.forward``
(``/root/reference/kws/libs/models.py:160-183``) and
``DepthwiseSeparableConvBlock.forward`` (``:108-119``) as a pure function of a
``state_dict`` with the reference's key names (``:135-144``):

    conv1.{weight[C,1,10,10],bias[C]}
    dsconv{i}.depthwise.{weight[C,1,3,3],bias[C]}
    dsconv{i}.pointwise.{weight[C,C,1,1],bias[C]}       i = 1..num_blocks
    fc.{weight[num_classes,C],bias[num_classes]}

The reference has C=64 and four blocks; ``channels``/``num_blocks`` are inferred
from the state_dict so the same function is the oracle for the "DS-CNN-L"
extension (SURVEY.md a-X1), which only changes those two numbers and keeps every
quirk: no activation between depthwise and pointwise, and the pointwise conv
carrying ``padding=1`` (``models.py:104-106``) so each block grows H and W by 2
with border pixels equal to ``ReLU(bias)``.

Pinned against the real reference class by ``tests/golden/dscnn_*.npz``
(``tests/golden/make_golden.py`` imports ``/root/reference/kws/libs/models.py``).
"""

from __future__ import annotations

import math
from typing import Dict

import torch
import torch.nn.functional as F


def num_blocks_of(sd: Dict[str, torch.Tensor]) -> int:
    n = 0
    while f"dsconv{n + 1}.depthwise.weight" in sd:
        n += 1
    return n


def forward(sd: Dict[str, torch.Tensor], x: torch.Tensor, return_layers: bool = False):
    """Logits ``[B, num_classes]`` for features ``x[B,1,T,F]`` (no softmax)."""
    layers = []
    c = sd["conv1.weight"].shape[0]
    y = F.relu(F.conv2d(x, sd["conv1.weight"], sd["conv1.bias"], stride=2, padding=2))
    layers.append(y)
    for i in range(1, num_blocks_of(sd) + 1):
        y = F.conv2d(y, sd[f"dsconv{i}.depthwise.weight"], sd[f"dsconv{i}.depthwise.bias"],
                     stride=1, padding=1, groups=c)
        y = F.conv2d(y, sd[f"dsconv{i}.pointwise.weight"], sd[f"dsconv{i}.pointwise.bias"],
                     stride=1, padding=1)
        y = F.relu(y)
        layers.append(y)
    y = F.adaptive_avg_pool2d(y, (1, 1)).view(y.size(0), -1)
    layers.append(y)
    logits = F.linear(y, sd["fc.weight"], sd["fc.bias"])
    return (logits, layers) if return_layers else logits


def init_state_dict(num_classes: int = 12, channels: int = 64, num_blocks: int = 4,
                    seed: int = 0, bias_std: float = 0.1, dtype=torch.float32):
    """Synthetic weights: the reference initialiser
    (``/root/reference/kws/libs/models.py:149-158``: conv ``kaiming_normal_(fan_out,
    relu)``, linear ``normal_(0, .01)``, biases 0) drawn from a seeded generator, then
    every bias overwritten with ``N(0, bias_std)`` so the padded-pointwise border rule
    is exercised."""
    g = torch.Generator().manual_seed(seed)

    def kaiming(shape):
        fan_out = shape[0] * shape[2] * shape[3]
        # torch's fan_out for conv weights is out_channels * kh * kw (groups ignored)
        std = math.sqrt(2.0 / fan_out)
        return torch.randn(shape, generator=g, dtype=torch.float64) * std

    def bias(n):
        return torch.randn(n, generator=g, dtype=torch.float64) * bias_std

    sd = {"conv1.weight": kaiming((channels, 1, 10, 10)), "conv1.bias": bias(channels)}
    for i in range(1, num_blocks + 1):
        sd[f"dsconv{i}.depthwise.weight"] = kaiming((channels, 1, 3, 3))
        sd[f"dsconv{i}.depthwise.bias"] = bias(channels)
        sd[f"dsconv{i}.pointwise.weight"] = kaiming((channels, channels, 1, 1))
        sd[f"dsconv{i}.pointwise.bias"] = bias(channels)
    sd["fc.weight"] = torch.randn((num_classes, channels), generator=g, dtype=torch.float64) * 0.01
    sd["fc.bias"] = bias(num_classes)
    return {k: v.to(dtype) for k, v in sd.items()}


def output_hw(T: int, Fdim: int, num_blocks: int = 4):
    """Spatial size after conv1 (k10 s2 p2) and ``num_blocks`` growing blocks."""
    h = (T + 4 - 10) // 2 + 1
    w = (Fdim + 4 - 10) // 2 + 1
    return h + 2 * num_blocks, w + 2 * num_blocks


def cross_entropy_train_step(sd, x, labels, state=None, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
    """Oracle for one training step (SURVEY.md a-T): ``CrossEntropyLoss()`` (mean) on
    the logits, backward, ``Adam(lr=1e-3)`` defaults
    (``/root/reference/kws/libs/training.py:112-123,285-297``).  Returns
    ``(loss, grads, new_sd, new_state)``; pure autograd on CPU."""
    params = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    logits = forward(params, x)
    loss = F.cross_entropy(logits, labels)
    loss.backward()
    grads = {k: p.grad.detach().clone() for k, p in params.items()}
    if state is None:
        state = {"step": 0, "m": {k: torch.zeros_like(v) for k, v in sd.items()},
                 "v": {k: torch.zeros_like(v) for k, v in sd.items()}}
    step = state["step"] + 1
    new_sd, m_new, v_new = {}, {}, {}
    for k, p in sd.items():
        g = grads[k]
        m = betas[0] * state["m"][k] + (1 - betas[0]) * g
        v = betas[1] * state["v"][k] + (1 - betas[1]) * g * g
        mhat = m / (1 - betas[0] ** step)
        vhat = v / (1 - betas[1] ** step)
        new_sd[k] = p - lr * mhat / (vhat.sqrt() + eps)
        m_new[k], v_new[k] = m, v
    return loss.detach(), grads, new_sd, {"step": step, "m": m_new, "v": v_new}
