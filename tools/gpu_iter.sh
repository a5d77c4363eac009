#!/bin/bash
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ransac_solve -s 12 -c 1 -o gpurun_out/k3_full python bench.py --steps 1 --warmup 3 --no-cpu --no-configs > gpurun_out/ncu_k3.log 2>&1
ncu -i gpurun_out/k3_full.ncu-rep --page source --csv > gpurun_out/k3_source.csv 2>/dev/null
ls -la gpurun_out/k3*
