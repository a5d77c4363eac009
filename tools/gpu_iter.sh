#!/bin/bash
# iteration call: K6 correctness (+ variants), K1 timing through the bench, GPU test suite
mkdir -p gpurun_out
timeout 300 python tools/gpu_check_l2.py > gpurun_out/l2_check.log 2>&1; echo "rc=$?" >> gpurun_out/l2_check.log
timeout 300 python tools/gpu_check_l2.py 8 > gpurun_out/l2_check_f8.log 2>&1; echo "rc=$?" >> gpurun_out/l2_check_f8.log
timeout 300 python tools/gpu_l2_variants.py > gpurun_out/l2_variants.log 2>&1; echo "rc=$?" >> gpurun_out/l2_variants.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "rc=$?" >> gpurun_out/bench.err
