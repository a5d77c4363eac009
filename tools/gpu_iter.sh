#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/gpu_check_l2.py 128 2>&1 | grep 'ALL\|FAIL\|MISMATCH\|"N": 2000\|"B": 16' | cut -c1-190 > gpurun_out/l2_bf16_nodbg.log
for m in 0 1; do for f in 0 128; do timeout 120 python tools/gpu_l2_one.py $m $f; done; done >> gpurun_out/l2_bf16_nodbg.log 2>&1
timeout 300 python -m pytest tests/test_gpu_match_l2.py -x -q 2>&1 | tail -2 >> gpurun_out/l2_bf16_nodbg.log
cat gpurun_out/l2_bf16_nodbg.log
