#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_train.py tests/test_gpu_optim.py -m gpu -q -s > gpurun_out/pytest_train.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_train.log
grep -E "EResNet\(|passed|failed|FAILED|rc=" gpurun_out/pytest_train.log | cut -c1-330 | head -20
