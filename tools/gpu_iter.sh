#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "rc=$?" >> gpurun_out/bench.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:hamming_top2 -s 3 -c 1 -o gpurun_out/k1_full python bench.py --steps 1 --warmup 3 --no-cpu > gpurun_out/ncu_k1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:l2_top2_tc -s 4 -c 2 -o gpurun_out/k6_full python tools/gpu_check_l2.py > gpurun_out/ncu_k6.log 2>&1
