#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_nets.py -m gpu -x -q -s > gpurun_out/pytest_nets.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_nets.log
grep -E "Denoise|passed|failed|FAILED|Error|rc=" gpurun_out/pytest_nets.log | cut -c1-300 | head -20
SRB200_LIB=image_super_resolution_b200/libsrb200_dbg.so timeout 300 python tools/gpu_ubench_mix.py > gpurun_out/ubench_mix_r2.log 2>&1; echo "ubench rc=$?"; tail -30 gpurun_out/ubench_mix_r2.log
