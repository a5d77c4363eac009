#!/bin/bash
# One GPU-box call: parity tests, smoke, both bench arms, ncu launch lists (bench step, one bundle
# adjustment, one SLAM sequence) + full captures (K1, K6).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > gpurun_out/gpu.txt
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
timeout 600 python tools/gpu_check_l2.py > gpurun_out/l2_check.log 2>&1; echo "l2 rc=$?" >> gpurun_out/l2_check.log
timeout 600 python tests/tools/gpu_ba_profile.py > gpurun_out/ba_profile.log 2>&1
timeout 600 python tools/gpu_slam_traj.py orb 20 2000 5 > gpurun_out/slam_traj.log 2>&1
timeout 600 python tools/gpu_mapops_bw.py > gpurun_out/mapops_bw.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench_c2.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_launch.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_ba.csv python tools/gpu_ba_one.py 18 5000 16 5 > gpurun_out/ncu_ba.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_slam.csv python tools/gpu_slam_traj.py orb 5 2000 5 > gpurun_out/ncu_slam.log 2>&1
timeout 600 python tools/gpu_orb_bench.py 32 2000 10 > gpurun_out/orb_bench.json 2> gpurun_out/orb_bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_orb.csv python tools/gpu_orb_bench.py 32 2000 1 > gpurun_out/ncu_orb.log 2>&1
if [ "$1" = "full" ]; then
timeout 900 ncu --set full --clock-control none --import-source on -k regex:orb_fast -s 8 -c 1 -o gpurun_out/orb_fast_full python tools/gpu_orb_bench.py 32 2000 1 > gpurun_out/ncu_orb_fast.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:orb_describe -s 1 -c 1 -o gpurun_out/orb_describe_full python tools/gpu_orb_bench.py 32 2000 1 > gpurun_out/ncu_orb_describe.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:hamming_top2 -s 3 -c 1 -o gpurun_out/k1_full python bench.py --steps 1 --warmup 3 --no-cpu > gpurun_out/ncu_k1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:l2_top2_tc -c 1 -o gpurun_out/k6_full python tools/gpu_l2_variants.py > gpurun_out/ncu_k6.log 2>&1
fi
ls -la gpurun_out
