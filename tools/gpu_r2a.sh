#!/bin/bash
# round 2, GPU session A: tests, the new bench line, both arms, A/B of the slot stride, ncu of the fused kernel
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r2a_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r2a_pytest.log
python bench.py --steps 10 --warmup 3 > $O/r2a_bench.json 2> $O/r2a_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 5 --warmup 1 > $O/r2a_ref.json 2> $O/r2a_ref.err; echo "ref rc=$?"
for lib in libkws_b200.so libkws_b200_pad64.so; do
  echo "== $lib" >> $O/r2a_ab.txt
  KWS_LIB=$lib python tools/bench_model.py 16384 40 5 auto >> $O/r2a_ab.txt 2>&1
  KWS_LIB=$lib python tools/bench_model.py 16384 10 5 auto >> $O/r2a_ab.txt 2>&1
done
python tools/bench_model.py 2048 40 1 auto > $O/r2a_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:dscnn_fused -s 2 -c 1 -o $O/r2a_fused python tools/bench_model.py 2048 40 1 auto > $O/r2a_ncu_fused.log 2>&1
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra > $O/r2a_plain2.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2a_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra > $O/r2a_ncu_launch.log 2>&1
ls -la $O | tail -20
