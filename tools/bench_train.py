"""Training-step throughput (BASELINE config 5): samples/s of forward+backward+allreduce+Adam on synthetic
Speech-Commands-shaped features, one process per GPU (torchrun for N>1), per-GPU batch fixed (weak scaling).
usage: [torchrun ...] bench_train.py [per_gpu_batch] [F] [steps]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import kws_b200  # noqa: F401,E402
from kws_b200.models import DepthwiseSeparableConv  # noqa: E402
from kws_b200.training import KWSTrainStep, TrainingConfig  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1028
F = int(sys.argv[2]) if len(sys.argv) > 2 else 10
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 10
rank, local, world = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("LOCAL_RANK", 0), ("WORLD_SIZE", 1)))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
torch.manual_seed(0)
model = DepthwiseSeparableConv()
ts = KWSTrainStep(model, TrainingConfig(batch_size=B), device=dev)
g = torch.Generator(device=dev).manual_seed(rank)
x = torch.randn(B, 1, 99, F, device=dev, generator=g)
y = torch.randint(0, 12, (B,), device=dev, generator=g)


def sync():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


for _ in range(3):
    ts.step(x, y)
sync()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
t_fb = t_ar = t_opt = 0.0
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(steps):
    ev[0].record(); ts.forward_backward(x, y)
    ev[1].record(); ts.allreduce_gradients()
    ev[2].record(); ts.optimizer_step()
    ev[3].record()
e1.record()
sync()
total_ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
if world > 1:
    dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
# per-phase split of the last step (events of earlier steps were overwritten)
t_fb, t_ar, t_opt = ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]), ev[2].elapsed_time(ev[3])
if rank == 0:
    ms = float(total_ms.item()) / steps
    print(json.dumps({"metric": "training samples/sec (DS-CNN reference net, 12 classes, fwd+bwd+allreduce+Adam)",
                      "value": world * B / (ms * 1e-3), "unit": "samples/s", "n_gpus": world, "ms_per_step": ms,
                      "per_gpu_batch": B, "features": [99, F], "last_step_ms": {"fwd_bwd": t_fb, "allreduce": t_ar, "adam": t_opt},
                      "allreduce_bytes": ts.grads.numel() * 4, "loss": float(ts.loss.item())}))
if world > 1:
    dist.destroy_process_group()
