#!/bin/bash
# round 2, profile session: full GPU suite, smoke, both bench arms, ncu launch list and full captures
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q > $O/r2p_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/r2p_pytest.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $O/r2p_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/r2p_smoke.log
timeout 600 python bench.py --impl reference --steps 10 --warmup 2 > $O/r2p_ref.json 2> $O/r2p_ref.err; echo "ref rc=$?"
timeout 900 python bench.py --steps 20 --warmup 5 > $O/r2p_bench.json 2> $O/r2p_bench.err; echo "bench rc=$?"
timeout 600 python tools/mfcc_tail.py 16384 uniform > $O/r2p_tail.txt 2>&1
timeout 300 python tools/bench_single.py > $O/r2p_split.txt 2>&1
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra > $O/r2p_plain2.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2p_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra > $O/r2p_ncu_launch.log 2>&1
python tools/run_single.py 65536 > $O/r2p_plain3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:pcm_logits_fused -s 2 -c 1 -o $O/r2p_single python tools/run_single.py 65536 > $O/r2p_ncu_single.log 2>&1
python tools/time_mfcc.py 8192 2 > $O/r2p_plain4.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:mfcc512_ws -s 3 -c 1 -o $O/r2p_mfcc python tools/time_mfcc.py 8192 2 > $O/r2p_ncu_mfcc.log 2>&1
tail -n 3 $O/r2p_pytest.log $O/r2p_smoke.log; cat $O/r2p_tail.txt; head -c 700 $O/r2p_bench.json
