#!/bin/bash
# round 2, GPU session B: single-launch PCM -> logits + float64 low-bin warp (tests first, under timeouts), then timing
O=gpurun_out; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_features.py -x -q > $O/r2b_features.log 2>&1; echo "features rc=$?" | tee -a $O/r2b_features.log
timeout 300 python -m pytest tests/test_gpu_single_launch.py -x -q > $O/r2b_single.log 2>&1; echo "single rc=$?" | tee -a $O/r2b_single.log
timeout 900 python -m pytest tests -m gpu -q > $O/r2b_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/r2b_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $O/r2b_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/r2b_smoke.log
for lb in 1 0; do echo "== KWS_LOW_BINS=$lb" >> $O/r2b_tail.txt; KWS_LOW_BINS=$lb timeout 300 python tools/mfcc_tail.py 8192 uniform >> $O/r2b_tail.txt 2>&1
  KWS_LOW_BINS=$lb timeout 100 python tools/time_mfcc.py >> $O/r2b_tail.txt 2>&1; done
timeout 300 python tools/bench_single.py > $O/r2b_split.txt 2>&1; echo "split rc=$?"
timeout 600 python bench.py --steps 10 --warmup 3 > $O/r2b_bench.json 2> $O/r2b_bench.err; echo "bench rc=$?"
tail -4 $O/r2b_features.log $O/r2b_single.log $O/r2b_pytest.log $O/r2b_smoke.log; cat $O/r2b_tail.txt $O/r2b_split.txt
