#!/bin/bash
# K1f (fp4 Hamming) bring-up: probe, parity against the popc kernel, timing; then the GPU suite
mkdir -p gpurun_out
timeout 120 python tools/gpu_check_hamming_f4.py probe > gpurun_out/f4_probe.log 2>&1; echo "rc=$?" >> gpurun_out/f4_probe.log
timeout 300 python tools/gpu_check_hamming_f4.py check > gpurun_out/f4_check.log 2>&1; echo "rc=$?" >> gpurun_out/f4_check.log
timeout 300 python tools/gpu_check_hamming_f4.py time > gpurun_out/f4_time.log 2>&1; echo "rc=$?" >> gpurun_out/f4_time.log
if [ "$1" = "suite" ]; then
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
fi
cat gpurun_out/f4_probe.log; cat gpurun_out/f4_check.log; cut -c1-230 gpurun_out/f4_time.log
