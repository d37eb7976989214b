"""Worker of tests/test_gpu_training.py::test_two_rank_nccl_step_equals_single_process (run under torchrun):
each rank trains on its (unequal) shard, gradients meet in ONE NCCL all-reduce per step; rank 0 also runs the
single-process step on the whole batch and compares."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import kws_b200  # noqa: F401,E402
from kws_b200.models import DepthwiseSeparableConv  # noqa: E402
from kws_b200.training import KWSTrainStep, TrainingConfig  # noqa: E402
from oracle import dscnn_ref  # noqa: E402

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
sd = dscnn_ref.init_state_dict(seed=9)
g = torch.Generator().manual_seed(10)
B = 16
x = torch.randn(B, 1, 99, 10, generator=g)
y = torch.randint(0, 12, (B,), generator=g)
m = DepthwiseSeparableConv()
m.load_state_dict(sd)
ts = KWSTrainStep(m, TrainingConfig(), device=dev)          # built BEFORE the process group exists
dist.init_process_group("nccl", device_id=dev)
assert ts.world == world
lo, hi = (0, 9) if rank == 0 else (9, B)
for _ in range(3):
    ts.step(x[lo:hi], y[lo:hi], global_batch=B)
mine = ts.params.clone()
other = mine.clone()
dist.broadcast(other, src=1)
ranks_equal = bool(torch.equal(mine, other)) if rank == 0 else True
if rank == 0:
    m1 = DepthwiseSeparableConv()
    m1.load_state_dict(sd)

    class _Solo(KWSTrainStep):
        world = 1
    solo = _Solo(m1, TrainingConfig(), device=dev)
    for _ in range(3):
        solo.step(x, y)
    d = (solo.params - mine).abs()
    # gradient parity between the sharded and the whole-batch step (same summation, different order)
    gd = float((solo.grads - ts.grads / world).abs().max().item() / solo.grads.abs().max().item())   # the reduced buffer is the SUM; Adam applies 1/world
    json.dump({"world": world, "max_abs_diff_vs_single": float(d.max().item()), "frac_within_2e-5": float((d < 2e-5).float().mean().item()),
               "grad_rel_diff": gd, "ranks_equal": ranks_equal}, open(sys.argv[1], "w"))
dist.barrier()
dist.destroy_process_group()
