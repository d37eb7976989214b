#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_features.py tests/test_gpu_single_launch.py tests/test_gpu_augment.py -x -q > $O/r2c_tests.log 2>&1; echo "tests rc=$?"
for lb in 1 0; do echo "== KWS_LOW_BINS=$lb" >> $O/r2c_time.txt; KWS_LOW_BINS=$lb timeout 100 python tools/time_mfcc.py >> $O/r2c_time.txt 2>&1; done
timeout 300 python tools/bench_single.py > $O/r2c_split.txt 2>&1
tail -n 3 $O/r2c_tests.log; cat $O/r2c_time.txt $O/r2c_split.txt
