"""Training step of the reference KWS network on sm_100a kernels.

Mirrors the step body of ``KWSTrainer.train_epoch`` (reference ``kws/libs/training.py:285-303``;
duplicate ``train.py:39-54``)::

    optimizer.zero_grad(); outputs = model(inputs); loss = criterion(outputs, labels)
    loss.backward(); optimizer.step()

with ``criterion = nn.CrossEntropyLoss()`` (``:115``) and ``optimizer = Adam(lr, weight_decay=0)``
(``:119-123``).  ``TrainingConfig`` keeps the reference's field names (``:25-66``); only the fields
the step itself uses are honoured (``learning_rate``, ``batch_size``, ``weight_decay`` must be 0,
``optimizer`` must be ``"adam"``).  The epoch loop, checkpoints, ClearML and tqdm of the reference
trainer are host-side control plane and out of scope.

Data parallelism (BASELINE config 5): one process per GPU, the batch is sharded by the caller, and
the only collective is ONE ``all_reduce(SUM)`` over the flat 26 444-float gradient buffer between
``kws_train_forward_backward`` and ``kws_train_adam``; the ``1/world_size`` mean is folded into the
Adam kernel (``grad_scale``).  No CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Any, Dict, Optional, Tuple

import torch

from . import _native
from .common.errors import ModelError
from .models import DepthwiseSeparableConv


@dataclass
class TrainingConfig:
    """Reference ``TrainingConfig`` (training.py:25-66), same names and defaults."""
    epochs: int = 10
    learning_rate: float = 0.001
    batch_size: int = 64
    num_workers: int = 4
    pin_memory: bool = True
    shuffle: bool = True
    weight_decay: float = 0.0
    optimizer: str = "adam"
    scheduler: Optional[str] = None
    scheduler_params: Dict[str, Any] = field(default_factory=dict)
    checkpoint_dir: str = "checkpoints"
    log_interval: int = 10
    use_clearml: bool = False
    project_name: str = "KeywordSpotting"

    def to_dict(self) -> Dict[str, Any]:
        return dict(self.__dict__)


def flatten_state_dict(sd: Dict[str, torch.Tensor], device) -> torch.Tensor:
    """state_dict order (models.py:135-144), torch layouts, one flat float32 buffer."""
    return torch.cat([v.detach().reshape(-1).to(device=device, dtype=torch.float32) for v in sd.values()]).contiguous()


class KWSTrainStep:
    """``step(inputs, labels) -> (loss, logits)``: forward, mean cross-entropy, backward, Adam.

    ``model`` is a :class:`DepthwiseSeparableConv` of the reference shape; its parameters are
    copied into a flat device buffer that the kernels update in place.  ``sync_to_model()`` writes
    them back into the module (and therefore into ``state_dict()`` / ``save()``).
    """

    def __init__(self, model: DepthwiseSeparableConv, config: Optional[TrainingConfig] = None,
                 device: Optional[torch.device] = None, process_group=None,
                 betas: Tuple[float, float] = (0.9, 0.999), eps: float = 1e-8):
        self.config = config if config is not None else TrainingConfig()
        if self.config.optimizer.lower() != "adam" or self.config.weight_decay != 0.0:
            raise ModelError("the sm_100a training step implements Adam with weight_decay=0 "
                             "(the reference's defaults, training.py:119-123)")
        if model.channels != 64 or model.num_blocks != 4 or model.num_classes > 32:
            raise ModelError("the training step covers the reference network only (64 channels, 4 blocks)")
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        if self.device.type != "cuda":
            raise ModelError("kws_b200 runs on CUDA devices only (no CPU fallback)")
        self._idx = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.model = model.to(self.device)
        # a native handle of its own: the step repacks the flat parameters into this handle's kernel
        # layouts every iteration, which must not disturb the weights the module's inference handle holds
        lib = _native.lib()
        self._h = _native.Handle(self._idx)
        _native.check(lib.kws_model_create(self._h.ptr, model.num_classes, model.channels, model.num_blocks), ModelError)
        for name, t in self.model.state_dict().items():
            a = t.detach().to("cpu", torch.float32).contiguous()
            shape = (C.c_int64 * a.dim())(*a.shape)
            _native.check(lib.kws_model_load_tensor(self._h.ptr, name.encode(), C.c_void_p(a.data_ptr()), shape, a.dim()),
                          ModelError)
        _native.check(lib.kws_model_finalize(self._h.ptr), ModelError)
        self._keys = list(self.model.state_dict().keys())
        self._shapes = [tuple(v.shape) for v in self.model.state_dict().values()]
        self.params = flatten_state_dict(self.model.state_dict(), self.device)
        n = int(_native.lib().kws_train_num_params(self._h.ptr))
        if n != self.params.numel():
            raise ModelError(f"flat parameter count mismatch: library {n}, state_dict {self.params.numel()}")
        self.grads = torch.zeros_like(self.params)
        self.exp_avg = torch.zeros_like(self.params)
        self.exp_avg_sq = torch.zeros_like(self.params)
        self.loss = torch.zeros(1, dtype=torch.float32, device=self.device)
        self.step_count = 0
        self.betas, self.eps = betas, eps
        self.pg = process_group
        self.world = 1
        if torch.distributed.is_available() and torch.distributed.is_initialized():
            self.world = torch.distributed.get_world_size(process_group)

    # -- the two kernels-side halves, exposed so the allreduce in between can be timed separately ----
    def forward_backward(self, inputs: torch.Tensor, labels: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        if inputs.dim() != 4 or inputs.shape[1] != 1:
            raise ModelError(f"expected input of shape (batch, 1, time, features), got {tuple(inputs.shape)}")
        x = inputs.to(self.device, torch.float32).contiguous()
        y = labels.to(self.device, torch.int64).contiguous()
        B, _, T, F = x.shape
        logits = torch.empty((B, self.model.num_classes), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            _native.check(_native.lib().kws_train_forward_backward(
                self._h.ptr, C.c_void_p(self.params.data_ptr()), C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()),
                B, T, F, C.c_void_p(self.grads.data_ptr()), C.c_void_p(self.loss.data_ptr()),
                C.c_void_p(logits.data_ptr()), _native.current_stream_ptr(self._idx)), ModelError)
        return self.loss, logits

    def allreduce_gradients(self) -> None:
        """The path's one collective: SUM over ranks of the flat gradient buffer (NCCL)."""
        if self.world > 1:
            torch.distributed.all_reduce(self.grads, op=torch.distributed.ReduceOp.SUM, group=self.pg)

    def optimizer_step(self) -> None:
        self.step_count += 1
        with torch.cuda.device(self.device):
            _native.check(_native.lib().kws_train_adam(
                self._h.ptr, C.c_void_p(self.params.data_ptr()), C.c_void_p(self.grads.data_ptr()),
                C.c_void_p(self.exp_avg.data_ptr()), C.c_void_p(self.exp_avg_sq.data_ptr()), self.params.numel(),
                self.step_count, float(self.config.learning_rate), float(self.betas[0]), float(self.betas[1]),
                float(self.eps), 1.0 / self.world, _native.current_stream_ptr(self._idx)), ModelError)

    def step(self, inputs: torch.Tensor, labels: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        """One optimisation step; returns ``(loss[1] device tensor of this rank's shard, logits)``.
        Unlike the reference (``loss.item()`` every step, training.py:300) nothing is synchronised."""
        loss, logits = self.forward_backward(inputs, labels)
        self.allreduce_gradients()
        self.optimizer_step()
        return loss, logits

    # -- views ------------------------------------------------------------------------------------
    def _split(self, flat: torch.Tensor) -> Dict[str, torch.Tensor]:
        out, o = {}, 0
        for k, shp in zip(self._keys, self._shapes):
            n = 1
            for d in shp:
                n *= d
            out[k] = flat[o:o + n].view(shp)
            o += n
        return out

    def state_dict(self) -> Dict[str, torch.Tensor]:
        return {k: v.clone() for k, v in self._split(self.params).items()}

    def gradients(self) -> Dict[str, torch.Tensor]:
        return {k: v.clone() for k, v in self._split(self.grads).items()}

    def sync_to_model(self) -> None:
        self.model.load_state_dict(self.state_dict())
