#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_train.py tests/test_gpu_optim.py -m gpu -x -q -s > gpurun_out/pytest_train.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_train.log
grep -E "EResNet\(|passed|failed|FAILED|Error|rc=" gpurun_out/pytest_train.log | cut -c1-300 | head -20
timeout 300 python tools/gpu_train_phases.py flat > gpurun_out/train_phases.log 2>&1; tail -14 gpurun_out/train_phases.log
