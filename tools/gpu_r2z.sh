#!/bin/bash
# round 2, closing session: full GPU suite, smoke, both bench arms, launch list, ncu captures of the DS-CNN-L kernels
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q > $O/r2z_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/r2z_pytest.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $O/r2z_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/r2z_smoke.log
timeout 600 python bench.py --impl reference --steps 10 --warmup 2 > $O/r2z_ref.json 2> $O/r2z_ref.err; echo "ref rc=$?"
timeout 900 python bench.py --steps 20 --warmup 5 > $O/r2z_bench.json 2> $O/r2z_bench.err; echo "bench rc=$?"
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra > $O/r2z_plain2.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2z_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra > $O/r2z_ncu_launch.log 2>&1
python tools/bench_dscnn_l.py 2048 1 > $O/r2z_l.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file $O/r2z_l_launches.csv python tools/bench_dscnn_l.py 2048 1 > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:dscnn_block_ws -s 7 -c 1 -o $O/r2z_block_ws python tools/bench_dscnn_l.py 2048 1 > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:dscnn_conv1 -s 2 -c 1 -o $O/r2z_conv1 python tools/bench_dscnn_l.py 2048 1 > /dev/null 2>&1
tail -n 3 $O/r2z_pytest.log $O/r2z_smoke.log; cat $O/r2z_l.log; head -c 600 $O/r2z_bench.json
