#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pnp.py tests/test_gpu_config_c5.py tests/test_gpu_tracks.py tests/test_compat_reference_drivers.py -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python tools/gpu_pnp_profile.py > gpurun_out/pnp_time.log 2>&1
timeout 300 python tools/gpu_check_pnp.py > gpurun_out/pnp_check.log 2>&1; echo "rc=$?" >> gpurun_out/pnp_check.log
