#!/bin/bash
# one GPU-box call while iterating (K6i timing variants + capture)
mkdir -p gpurun_out
timeout 300 python tools/gpu_sanitize_cases.py > gpurun_out/small_cases.log 2>&1; echo "rc=$?" >> gpurun_out/small_cases.log
timeout 300 python tools/gpu_check_l2.py 256 2>&1 | grep '"mode"' > gpurun_out/l2_i8_norank.log
timeout 300 python tools/gpu_check_l2.py 768 2>&1 | grep '"mode"' > gpurun_out/l2_i8_noload.log
timeout 300 python tools/gpu_check_l2.py 2>&1 | grep '"mode"\|ALL' > gpurun_out/l2_i8_default.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:match_kernel -s 40 -c 1 -o gpurun_out/k6i_full python tools/gpu_check_l2.py > gpurun_out/ncu_k6i.log 2>&1
tail -3 gpurun_out/small_cases.log; cat gpurun_out/l2_i8_norank.log gpurun_out/l2_i8_noload.log gpurun_out/l2_i8_default.log | cut -c1-200
