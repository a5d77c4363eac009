#!/bin/bash
# after K1f became the default tensor path: its checks, the GPU suite, smoke, bench
mkdir -p gpurun_out
timeout 300 python tools/gpu_check_hamming_f4.py all > gpurun_out/f4_all.log 2>&1; echo "rc=$?" >> gpurun_out/f4_all.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
grep -v '"path": "popc"' gpurun_out/f4_all.log | cut -c1-160; tail -3 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log; tail -2 gpurun_out/bench.err; head -c 1500 gpurun_out/bench.json
