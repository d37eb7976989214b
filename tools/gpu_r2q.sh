#!/bin/bash
# round 2, session q: full GPU suite + smoke + bench on the current head, ncu of the DS-CNN-L kernel and the training step
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q > $O/r2q_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/r2q_pytest.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $O/r2q_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/r2q_smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > $O/r2q_bench.json 2> $O/r2q_bench.err; echo "bench rc=$?"
python tools/bench_dscnn_l.py 2048 2 > $O/r2q_l.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:dscnn_gemm_kernel -s 15 -c 1 -o $O/r2q_gemm python tools/bench_dscnn_l.py 2048 1 > $O/r2q_ncu_gemm.log 2>&1
python tools/bench_train.py 1028 10 5 > $O/r2q_train.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $O/r2q_train_launches.csv python tools/bench_train.py 1028 10 2 > $O/r2q_ncu_train.log 2>&1
tail -n 3 $O/r2q_pytest.log $O/r2q_smoke.log; cat $O/r2q_l.log $O/r2q_train.log; head -c 400 $O/r2q_bench.json
