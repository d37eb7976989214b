#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_model.py tests/test_gpu_single_launch.py tests/test_gpu_streaming.py tests/test_gpu_training.py -x -q > $O/r2d_tests.log 2>&1; echo "tests rc=$?"
timeout 200 python tools/bench_model.py 16384 40 5 auto > $O/r2d_model.txt 2>&1
timeout 200 python tools/bench_model.py 16384 10 5 auto >> $O/r2d_model.txt 2>&1
python tools/bench_model.py 2048 40 1 auto > $O/r2d_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:dscnn_fused -s 2 -c 1 -o $O/r2d_fused python tools/bench_model.py 2048 40 1 auto > $O/r2d_ncu.log 2>&1
tail -n 4 $O/r2d_tests.log; cat $O/r2d_model.txt
