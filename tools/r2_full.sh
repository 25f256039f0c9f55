#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 300 python tools/gpu_train_phases.py flat > gpurun_out/train_phases.log 2>&1; tail -3 gpurun_out/train_phases.log
