#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_match_l2.py tests/test_gpu_pose.py -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python tools/gpu_check_l2.py > gpurun_out/l2_check.log 2>&1; echo "rc=$?" >> gpurun_out/l2_check.log
timeout 300 python tools/gpu_check_l2.py 16 > gpurun_out/l2_check_f16.log 2>&1; echo "rc=$?" >> gpurun_out/l2_check_f16.log
timeout 300 python tools/gpu_check_l2.py 32 > gpurun_out/l2_check_f32.log 2>&1; echo "rc=$?" >> gpurun_out/l2_check_f32.log
