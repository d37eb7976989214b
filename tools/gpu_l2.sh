#!/bin/bash
# DS-CNN-L: launch list of one forward and a full ncu capture of one warp-specialised block launch
O=gpurun_out; mkdir -p $O
python tools/bench_dscnn_l.py 2048 1 > $O/l2_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/l2_launches.csv python tools/bench_dscnn_l.py 2048 1 > $O/l2_ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:dscnn_block_ws -s 7 -c 1 -o $O/l2_ws python tools/bench_dscnn_l.py 2048 1 > $O/l2_ncu_ws.log 2>&1
cat $O/l2_plain.log
