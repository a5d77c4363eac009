#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/gpu_check_l2.py 2>&1 | grep 'ALL\|FAIL\|MISMATCH' > gpurun_out/l2_i8_ok.log
timeout 300 python tools/gpu_check_l2.py 256 2>&1 | grep 'ALL\|FAIL\|MISMATCH' >> gpurun_out/l2_i8_ok.log
timeout 300 python tools/gpu_check_l2.py 512 2>&1 | grep 'ALL\|FAIL\|MISMATCH' >> gpurun_out/l2_i8_ok.log
for m in 0 1; do for f in 0 256 512 128; do timeout 120 python tools/gpu_l2_one.py $m $f; done; done > gpurun_out/l2_one.log 2>&1
for f in 2 10 26 1; do timeout 120 python tools/gpu_hamming_one.py $f; done > gpurun_out/k1t_one.log 2>&1
timeout 300 python -m pytest tests/test_gpu_match.py tests/test_gpu_match_l2.py -x -q 2>&1 | tail -2 > gpurun_out/pytest_match.log
cat gpurun_out/l2_i8_ok.log gpurun_out/l2_one.log gpurun_out/k1t_one.log gpurun_out/pytest_match.log | cut -c1-200
