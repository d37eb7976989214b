#!/bin/bash
# A/B timing of library variants built with build.py --out=libkws_b200_<x>.so: DS-CNN-L forward at 2048 clips
O=gpurun_out; mkdir -p $O; : > $O/ab.log
for L in libkws_b200.so "$@"; do
  [ -f keyword-spotting_b200/$L ] || continue
  echo "== $L" | tee -a $O/ab.log
  KWS_LIB=$L timeout 120 python tools/bench_dscnn_l.py 2048 5 2>&1 | tail -n 1 | cut -c1-200 | tee -a $O/ab.log
done
