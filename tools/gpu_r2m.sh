#!/bin/bash
# round 2, multi-GPU session (8 GPUs of one box): NCCL training-step test, host->device ceiling for 1..8 concurrent
# ranks, the bench line at N=2 and N=8
O=gpurun_out; mkdir -p $O
nvidia-smi topo -m > $O/r2m_topo.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 python -m pytest tests/test_gpu_training.py -q -k "nccl" > $O/r2m_nccl.log 2>&1; echo "nccl test rc=$?" | tee -a $O/r2m_nccl.log
: > $O/r2m_pcie.jsonl
timeout 120 python tools/pcie_probe.py >> $O/r2m_pcie.jsonl 2>$O/r2m_pcie.err
for n in 2 4 8; do timeout 180 $TR --nproc-per-node $n --master-port $((29600+n)) tools/pcie_probe.py >> $O/r2m_pcie.jsonl 2>>$O/r2m_pcie.err; done
for n in 2 8; do timeout 400 $TR --nproc-per-node $n --master-port $((29700+n)) bench.py --gpus $n --steps 10 --warmup 3 > $O/r2m_bench_$n.json 2> $O/r2m_bench_$n.err; echo "bench $n rc=$?"; done
tail -n 3 $O/r2m_nccl.log; cat $O/r2m_pcie.jsonl; head -c 600 $O/r2m_bench_8.json
