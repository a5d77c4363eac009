#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_essential.py tests/test_gpu_pose.py tests/test_gpu_match_l2.py -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
timeout 900 python tools/gpu_configs.py > gpurun_out/configs.log 2>&1; echo rc=$? >> gpurun_out/configs.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "rc=$?" >> gpurun_out/bench.err
