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
                 betas: Tuple[float, float] = (0.9, 0.999), eps: float = 1e-8, check_labels: bool = False):
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
        self.check_labels = check_labels
        self._local_weight = 1.0     # B_local * world / B_global of the last forward_backward (1 for equal shards)

    @property
    def world(self) -> int:
        """Read at step time: a process group initialised after this object was built is honoured."""
        if torch.distributed.is_available() and torch.distributed.is_initialized():
            return torch.distributed.get_world_size(self.pg)
        return 1

    # -- the two kernels-side halves, exposed so the allreduce in between can be timed separately ----
    def forward_backward(self, inputs: torch.Tensor, labels: torch.Tensor,
                         global_batch: Optional[int] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        """Forward, mean cross-entropy over THIS rank's shard, backward into ``self.grads``.

        The kernels scale by ``1/B_local`` and the Adam kernel by ``1/world``, which is the global-batch mean
        the single-process step computes when every rank holds the same number of samples.  With unequal
        shards (``utils.sharding.shard_range`` hands out shards that differ by one clip) pass the summed batch
        size as ``global_batch``: the local gradient is then weighted by ``B_local * world / global_batch``
        before the all-reduce, so the result is again the exact global mean, with no extra collective."""
        if inputs.dim() != 4 or inputs.shape[1] != 1:
            raise ModelError(f"expected input of shape (batch, 1, time, features), got {tuple(inputs.shape)}")
        x = inputs.to(self.device, torch.float32).contiguous()
        y = labels.to(self.device, torch.int64).contiguous()
        if self.check_labels and y.numel() and (int(y.min()) < 0 or int(y.max()) >= self.model.num_classes):
            raise ModelError(f"label outside [0, {self.model.num_classes}) (torch's CrossEntropyLoss raises here too)")
        B, _, T, F = x.shape
        self._local_weight = 1.0 if global_batch is None else float(B) * self.world / float(global_batch)
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
            if self._local_weight != 1.0:
                self.grads.mul_(self._local_weight)
            torch.distributed.all_reduce(self.grads, op=torch.distributed.ReduceOp.SUM, group=self.pg)

    def optimizer_step(self) -> None:
        self.step_count += 1
        with torch.cuda.device(self.device):
            _native.check(_native.lib().kws_train_adam(
                self._h.ptr, C.c_void_p(self.params.data_ptr()), C.c_void_p(self.grads.data_ptr()),
                C.c_void_p(self.exp_avg.data_ptr()), C.c_void_p(self.exp_avg_sq.data_ptr()), self.params.numel(),
                self.step_count, float(self.config.learning_rate), float(self.betas[0]), float(self.betas[1]),
                float(self.eps), 1.0 / self.world, _native.current_stream_ptr(self._idx)), ModelError)

    def step(self, inputs: torch.Tensor, labels: torch.Tensor,
             global_batch: Optional[int] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        """One optimisation step; returns ``(loss[1] device tensor of this rank's shard, logits)``.
        Unlike the reference (``loss.item()`` every step, training.py:300) nothing is synchronised."""
        loss, logits = self.forward_backward(inputs, labels, global_batch)
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

    # -- torch.optim.Adam-compatible optimizer state (reference checkpoints, training.py:199-216,251-260) ----
    def optimizer_state_dict(self) -> Dict[str, Any]:
        """The layout ``torch.optim.Adam.state_dict()`` has for ``Adam(model.parameters())``: parameter index
        -> ``{step, exp_avg, exp_avg_sq}`` in ``model.parameters()`` order (= state_dict order here), one param
        group.  The reference's ``Trainer.load_checkpoint`` can load it into its own optimizer."""
        m, v = self._split(self.exp_avg), self._split(self.exp_avg_sq)
        state = {}
        if self.step_count > 0:
            for i, k in enumerate(self._keys):
                state[i] = {"step": torch.tensor(float(self.step_count)), "exp_avg": m[k].detach().cpu().clone(),
                            "exp_avg_sq": v[k].detach().cpu().clone()}
        group = {"lr": float(self.config.learning_rate), "betas": tuple(self.betas), "eps": float(self.eps),
                 "weight_decay": 0.0, "amsgrad": False, "maximize": False, "foreach": None, "capturable": False,
                 "differentiable": False, "fused": None, "decoupled_weight_decay": False,
                 "params": list(range(len(self._keys)))}
        return {"state": state, "param_groups": [group]}

    def load_optimizer_state_dict(self, osd: Dict[str, Any]) -> None:
        """Resume the Adam moments and the step count from a ``torch.optim.Adam`` state dict (e.g. the
        ``optimizer_state_dict`` of a checkpoint written by the reference trainer)."""
        groups = osd.get("param_groups", [])
        if len(groups) != 1 or len(groups[0].get("params", [])) != len(self._keys):
            raise ModelError("optimizer state does not match Adam(model.parameters()) of the reference network")
        if groups[0].get("weight_decay", 0.0) != 0.0 or groups[0].get("amsgrad", False):
            raise ModelError("only Adam with weight_decay=0, amsgrad=False is implemented (training.py:119-123)")
        self.betas = tuple(groups[0].get("betas", self.betas))
        self.eps = float(groups[0].get("eps", self.eps))
        self.config.learning_rate = float(groups[0].get("lr", self.config.learning_rate))
        st = osd.get("state", {})
        self.exp_avg.zero_(); self.exp_avg_sq.zero_()
        steps = set()
        m, v = self._split(self.exp_avg), self._split(self.exp_avg_sq)
        for i, k in enumerate(self._keys):
            ent = st.get(i, st.get(str(i)))
            if ent is None:
                continue
            m[k].copy_(ent["exp_avg"].to(self.device, torch.float32))
            v[k].copy_(ent["exp_avg_sq"].to(self.device, torch.float32))
            steps.add(int(float(ent["step"])))
        if len(steps) > 1:
            raise ModelError(f"parameters disagree on the Adam step count: {sorted(steps)}")
        self.step_count = steps.pop() if steps else 0

    def load_parameters(self, sd: Dict[str, torch.Tensor]) -> None:
        """Overwrite the flat training parameters from a state_dict (and the module's copy with them)."""
        self.params.copy_(flatten_state_dict({k: sd[k] for k in self._keys}, self.device))
        self.sync_to_model()


class KWSTrainer:
    """Drop-in for the reference ``KWSTrainer`` (kws/libs/training.py:69-260 ``Trainer`` + :263-393): same
    constructor, ``train_epoch`` / ``validate`` / ``train`` / ``save_checkpoint`` / ``load_checkpoint``, same
    ``metrics`` dict and checkpoint layout, with the step body (:285-297) replaced by
    :meth:`KWSTrainStep.step` and the validation forward by the inference kernels.

    Differences that are part of the contract: the device is always CUDA; ``loss.item()`` / the accuracy
    count still synchronise once per batch like the reference (``:300-303``) because ``train_epoch`` returns
    them; tqdm / ClearML / loguru decoration is not reproduced (host control plane).
    Checkpoints carry the Adam state in ``torch.optim.Adam``'s layout, so either trainer loads the other's files.
    """

    def __init__(self, model: DepthwiseSeparableConv, train_loader, val_loader, config: TrainingConfig,
                 device: Optional[torch.device] = None, process_group=None):
        from pathlib import Path
        self.model, self.train_loader, self.val_loader, self.config = model, train_loader, val_loader, config
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.model = self.model.to(self.device)
        self.metrics = {"train_loss": [], "train_acc": [], "val_loss": [], "val_acc": [], "best_val_acc": 0.0,
                        "best_epoch": 0}
        if config.scheduler:
            raise ModelError("learning-rate schedulers are host-side policy and not part of the step (out of scope)")
        self.step_fn = KWSTrainStep(self.model, config, device=self.device, process_group=process_group)
        Path(config.checkpoint_dir).mkdir(parents=True, exist_ok=True)

    def train_epoch(self, epoch: int) -> Tuple[float, float]:
        running_loss, correct, total, nb = 0.0, 0, 0, 0
        for inputs, labels in self.train_loader:
            labels = labels.to(self.device)
            loss, logits = self.step_fn.step(inputs, labels)
            running_loss += loss.item()                       # training.py:300
            correct += int((logits.argmax(1) == labels).sum().item())
            total += labels.size(0)
            nb += 1
        self.step_fn.sync_to_model()
        return running_loss / max(nb, 1), 100.0 * correct / max(total, 1)

    def validate(self, epoch: int) -> Tuple[float, float]:
        running_loss, correct, total, nb = 0.0, 0, 0, 0
        for inputs, labels in self.val_loader:
            labels = labels.to(self.device)
            logits = self.model(inputs.to(self.device))
            running_loss += float(torch.nn.functional.cross_entropy(logits, labels).item())
            correct += int((logits.argmax(1) == labels).sum().item())
            total += labels.size(0)
            nb += 1
        return running_loss / max(nb, 1), 100.0 * correct / max(total, 1)

    def train(self) -> Dict[str, Any]:
        for epoch in range(self.config.epochs):
            tl, ta = self.train_epoch(epoch)
            self.metrics["train_loss"].append(tl); self.metrics["train_acc"].append(ta)
            vl, va = self.validate(epoch)
            self.metrics["val_loss"].append(vl); self.metrics["val_acc"].append(va)
            if va > self.metrics["best_val_acc"]:
                self.metrics["best_val_acc"], self.metrics["best_epoch"] = va, epoch
                self.save_checkpoint("best_model.pth")
            if (epoch + 1) % 5 == 0:
                self.save_checkpoint(f"checkpoint_epoch_{epoch + 1}.pth")
        self.save_checkpoint("final_model.pth")
        return self.metrics

    def save_checkpoint(self, filename: str) -> str:
        from pathlib import Path
        path = Path(self.config.checkpoint_dir) / filename
        self.step_fn.sync_to_model()
        torch.save({"model_state_dict": {k: v.detach().cpu() for k, v in self.model.state_dict().items()},
                    "optimizer_state_dict": self.step_fn.optimizer_state_dict(), "metrics": self.metrics,
                    "config": self.config.to_dict(), "epoch": self.metrics.get("best_epoch", 0)}, path)
        return str(path)

    def load_checkpoint(self, checkpoint_path: str) -> None:
        ck = torch.load(checkpoint_path, map_location="cpu", weights_only=False)
        self.step_fn.load_parameters(ck["model_state_dict"])
        self.step_fn.load_optimizer_state_dict(ck["optimizer_state_dict"])
        self.metrics = ck.get("metrics", self.metrics)
