#!/bin/bash
# one GPU-box call: the new no-allocation entry point alone in a fresh process, then the whole suite
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/pytest_no_alloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_no_alloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_no_alloc_alone.log | cut -c1-200; tail -4 gpurun_out/pytest_gpu.log | cut -c1-200
