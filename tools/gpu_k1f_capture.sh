#!/bin/bash
# ncu evidence for K1f: the launch list of the bench step and one full capture of the matching kernel
mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu --no-configs > gpurun_out/bench_nocfg.json 2> gpurun_out/bench_nocfg.err || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench_c2.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-configs > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:match_kernel -s 3 -c 1 -o gpurun_out/k1f_full -f python bench.py --steps 1 --warmup 3 --no-cpu --no-configs > gpurun_out/ncu_k1f.log 2>&1
ls -la gpurun_out | head -30; tail -5 gpurun_out/ncu_k1f.log
