#!/bin/bash
# round 2, GPU session B: single-launch PCM -> logits (tests first, under a timeout), then timing and the role split
O=gpurun_out; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_single_launch.py -x -q > $O/r2b_single.log 2>&1; echo "single rc=$?" | tee -a $O/r2b_single.log
timeout 600 python -m pytest tests -m gpu -x -q > $O/r2b_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/r2b_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $O/r2b_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/r2b_smoke.log
timeout 300 python tools/bench_single.py > $O/r2b_split.txt 2>&1; echo "split rc=$?"
timeout 600 python bench.py --steps 10 --warmup 3 > $O/r2b_bench.json 2> $O/r2b_bench.err; echo "bench rc=$?"
tail -3 $O/r2b_single.log $O/r2b_pytest.log $O/r2b_smoke.log; cat $O/r2b_split.txt
