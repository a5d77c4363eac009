#!/bin/bash
# one GPU-box call while iterating: parity tests of the touched kernels, bench, launch list, K1t capture
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > gpurun_out/gpu.txt
timeout 900 python -m pytest tests/test_gpu_match.py tests/test_gpu_match_l2.py tests/test_gpu_pnp.py tests/test_gpu_pose.py tests/test_gpu_config_c5.py tests/test_gpu_essential.py tests/test_gpu_slam.py tests/test_gpu_device_map.py -x -q > gpurun_out/pytest_iter.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_iter.log
timeout 600 python tools/gpu_check_l2.py > gpurun_out/l2_check_int8.log 2>&1; echo "rc=$?" >> gpurun_out/l2_check_int8.log
timeout 600 python tools/gpu_check_l2.py 128 > gpurun_out/l2_check_bf16.log 2>&1; echo "rc=$?" >> gpurun_out/l2_check_bf16.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "rc=$?" >> gpurun_out/bench.err
if [ "$1" = "ncu" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench_c2.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-configs > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:match_kernel -s 3 -c 1 -o gpurun_out/k1t_full python bench.py --steps 1 --warmup 3 --no-cpu --no-configs > gpurun_out/ncu_k1t.log 2>&1
fi
tail -5 gpurun_out/pytest_iter.log; tail -8 gpurun_out/l2_check_int8.log; tail -8 gpurun_out/l2_check_bf16.log; tail -3 gpurun_out/bench.err; head -c 6000 gpurun_out/bench.json
