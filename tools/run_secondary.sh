#!/bin/bash
# End-of-round secondary benches on one B200: every line lands in gpurun_out/secondary.jsonl / secondary.txt
out=gpurun_out/secondary.jsonl; txt=gpurun_out/secondary.txt; : > $out; : > $txt
run() { echo "{\"cmd\": \"$*\"}" >> $out; timeout 300 "$@" 2>&1 | tail -1 >> $out; }
run python tools/bench_train.py 1028 10 10
run python tools/bench_train.py 1028 40 10
run python tools/bench_train.py 8192 10 5
run python tools/bench_crnn.py 16384 3
run python tools/bench_dscnn_l.py 1024 2
run python tools/bench_eager.py 16384 40
run python tools/bench_eager_l.py 1024
run python tools/bench_eager_crnn.py 16384
for F in 40 10; do echo "== python tools/bench_model.py 16384 $F 5" >> $txt; timeout 300 python tools/bench_model.py 16384 $F 5 2>&1 | tail -4 >> $txt; done
echo "== python tools/bench_model.py 256 40 20 auto (BASELINE configs[0] batch)" >> $txt; timeout 100 python tools/bench_model.py 256 40 20 auto 2>&1 | tail -3 >> $txt
