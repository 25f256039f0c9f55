#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_rs.py tests/test_gpu_configs.py -m gpu -q -x > gpurun_out/pytest_rs.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/pytest_rs.log
timeout 600 python tools/gpu_scale_rs.py 1 2>&1 | tee gpurun_out/scale_rs_1gpu.log | tail -12
timeout 300 python tools/gpu_image_bench.py 2>&1 | tee gpurun_out/image_bench.log | tail -8
