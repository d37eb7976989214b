#!/bin/bash
# DS-CNN-L iteration: parity tests of the tensor-core GEMM path, then timing with the warp-specialised block kernel on / off
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_model.py -q -x > $O/l_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/l_pytest.log
tail -n 4 $O/l_pytest.log
timeout 120 python tools/bench_dscnn_l.py 2048 3 2>&1 | tail -n 2 | tee $O/l_ws.log
KWS_GEMM_WS=0 timeout 120 python tools/bench_dscnn_l.py 2048 3 2>&1 | tail -n 1 | tee $O/l_old.log
timeout 120 python tools/bench_dscnn_l.py 8192 2 2>&1 | tail -n 1 | tee -a $O/l_ws.log
if [ -f keyword-spotting_b200/libkws_b200_prof.so ]; then KWS_LIB=libkws_b200_prof.so timeout 120 python tools/bench_dscnn_l.py 2048 1 2>&1 | grep -E "issuer|group 0" | tail -n 10 | tee $O/l_prof.log; fi
