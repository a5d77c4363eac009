#!/bin/bash
# one quick GPU-box call while iterating on K1t (tensor-core Hamming)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > gpurun_out/gpu.txt
timeout 600 python -m pytest tests/test_gpu_match.py -x -q > gpurun_out/pytest_match.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_match.log
timeout 300 python tools/gpu_check_hamming.py > gpurun_out/hamming_check.log 2>&1; echo "rc=$?" >> gpurun_out/hamming_check.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "rc=$?" >> gpurun_out/bench.err
tail -5 gpurun_out/pytest_match.log; tail -20 gpurun_out/hamming_check.log; cat gpurun_out/bench.json | head -c 1500
