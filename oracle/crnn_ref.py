"""CRNN / GRU model oracle (SURVEY.md a-X3): plain PyTorch on CPU.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  **Parity unpinned by the reference**: the
reference has no recurrent model (``grep -r 'GRU\\|LSTM\\|RNN' /root/reference`` is empty), so this is the
builder's own definition in the reference's idiom -- its conv1 (``kws/libs/models.py:135,170``:
``Conv2d(1, C, 10, stride=2, padding=2)`` + ReLU), the conv output read as a sequence over time,
``torch.nn.GRU`` semantics (gate order r, z, n; ``h_0 = 0``), and ``fc`` on the last hidden state
(``models.py:144,181``).  The GRU is written out with explicit matmuls so the oracle does not depend on
``nn.GRU``'s fused kernels; ``tests/test_oracle.py`` checks it against ``torch.nn.GRU`` itself."""
from __future__ import annotations

from typing import Dict

import torch
import torch.nn.functional as F


def forward(sd: Dict[str, torch.Tensor], x: torch.Tensor) -> torch.Tensor:
    """Logits for ``x[B,1,T,F]``; keys ``conv1.*``, ``gru.{weight,bias}_{ih,hh}_l{k}``, ``fc.*``."""
    y = F.relu(F.conv2d(x, sd["conv1.weight"], sd["conv1.bias"], stride=2, padding=2))   # [B, C, T', W']
    B, Cc, T1, W1 = y.shape
    seq = y.permute(0, 2, 1, 3).reshape(B, T1, Cc * W1)
    layer = 0
    while f"gru.weight_ih_l{layer}" in sd:
        wih, whh = sd[f"gru.weight_ih_l{layer}"], sd[f"gru.weight_hh_l{layer}"]
        bih, bhh = sd[f"gru.bias_ih_l{layer}"], sd[f"gru.bias_hh_l{layer}"]
        H = whh.shape[1]
        h = torch.zeros(B, H, dtype=x.dtype)
        outs = []
        for t in range(T1):
            gi = seq[:, t] @ wih.T + bih
            gh = h @ whh.T + bhh
            r = torch.sigmoid(gi[:, :H] + gh[:, :H])
            z = torch.sigmoid(gi[:, H:2 * H] + gh[:, H:2 * H])
            n = torch.tanh(gi[:, 2 * H:] + r * gh[:, 2 * H:])
            h = (1 - z) * n + z * h
            outs.append(h)
        seq = torch.stack(outs, 1)
        layer += 1
    return seq[:, -1] @ sd["fc.weight"].T + sd["fc.bias"]
