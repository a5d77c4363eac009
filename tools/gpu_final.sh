#!/bin/bash
# end-of-round check of the committed tree: GPU suite, smoke, both bench arms
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
tail -3 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log | cut -c1-160; tail -1 gpurun_out/bench.err; head -c 700 gpurun_out/bench.json; echo; head -c 300 gpurun_out/bench_ref.json
