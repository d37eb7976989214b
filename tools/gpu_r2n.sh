#!/bin/bash
# round 2, closing multi-GPU run (8 GPUs of one box): the bench line at N=8 on the final kernels
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 500 $TR --nproc-per-node 8 --master-port 29708 bench.py --gpus 8 --steps 10 --warmup 3 > $O/r2n_bench_8.json 2> $O/r2n_bench_8.err; echo "bench 8 rc=$?"
head -c 400 $O/r2n_bench_8.json; tail -n 3 $O/r2n_bench_8.err
