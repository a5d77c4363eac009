#!/bin/bash
# One GPU-box call: the whole GPU test suite, smoke, both bench arms, the per-kernel timing tools,
# the reference's own frame loop on the GPU path, compute-sanitizer on the hand-synchronised
# kernels (small cases), ncu launch list (+ full captures with "full").
mkdir -p gpurun_out
rm -f gpurun_out/parity.json
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > gpurun_out/gpu.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
timeout 600 python tools/gpu_check_hamming.py > gpurun_out/hamming_check.log 2>&1; echo "rc=$?" >> gpurun_out/hamming_check.log
timeout 600 python tools/gpu_check_hamming_f4.py all > gpurun_out/hamming_f4_check.log 2>&1; echo "rc=$?" >> gpurun_out/hamming_f4_check.log
timeout 600 python tools/gpu_check_l2.py > gpurun_out/l2_check_int8.log 2>&1; echo "rc=$?" >> gpurun_out/l2_check_int8.log
timeout 600 python tools/gpu_check_l2.py 128 > gpurun_out/l2_check_bf16.log 2>&1; echo "rc=$?" >> gpurun_out/l2_check_bf16.log
timeout 600 python tools/gpu_reference_loop.py > gpurun_out/reference_loop.log 2>&1; echo "rc=$?" >> gpurun_out/reference_loop.log
timeout 600 python tools/gpu_reference_loop_images.py 6 > gpurun_out/reference_loop_images.log 2>&1; echo "rc=$?" >> gpurun_out/reference_loop_images.log
timeout 600 python tools/gpu_check_sift.py > gpurun_out/sift_check.log 2>&1; echo "rc=$?" >> gpurun_out/sift_check.log
timeout 600 python tools/gpu_slam_traj.py orb 20 2000 5 > gpurun_out/slam_traj.log 2>&1
timeout 600 python tools/gpu_orb_bench.py 32 2000 10 > gpurun_out/orb_bench.json 2> gpurun_out/orb_bench.err
# compute-sanitizer is CLOSED on this pool ("runs under it have left GPUs needing a reset"), so the
# small cases of the hand-synchronised kernels run plain, each checked against cv2
timeout 300 python tools/gpu_sanitize_cases.py > gpurun_out/small_cases.log 2>&1; echo "rc=$?" >> gpurun_out/small_cases.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench_c2.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-configs > gpurun_out/ncu_launch.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_sift.csv python tools/gpu_sift_bench.py 480 640 32 1 > gpurun_out/ncu_launch_sift.log 2>&1
if [ "$1" = "full" ]; then
timeout 600 ncu --set full --clock-control none --import-source on -k regex:describe_kernel -c 1 -s 1 -o gpurun_out/sift_describe_full -f python tools/gpu_sift_bench.py 480 640 32 1 > gpurun_out/ncu_sift_describe.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:blur_kernel -c 1 -s 46 -o gpurun_out/sift_blur_full -f python tools/gpu_sift_bench.py 480 640 32 1 > gpurun_out/ncu_sift_blur.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:match_kernel -s 3 -c 1 -o gpurun_out/k1f_full -f python bench.py --steps 1 --warmup 3 --no-cpu --no-configs > gpurun_out/ncu_k1f.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:match_kernel -s 40 -c 1 -o gpurun_out/k6i_full -f python tools/gpu_check_l2.py > gpurun_out/ncu_k6i.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ransac_score -s 12 -c 1 -o gpurun_out/k2_full -f python bench.py --steps 1 --warmup 3 --no-cpu --no-configs > gpurun_out/ncu_k2.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ransac_solve -s 12 -c 1 -o gpurun_out/k3_full -f python bench.py --steps 1 --warmup 3 --no-cpu --no-configs > gpurun_out/ncu_k3.log 2>&1
fi
ls -la gpurun_out | head -60
tail -3 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log; head -c 3000 gpurun_out/bench.json
