#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_train.py -m gpu -x -q -k "resnet_bn or reference_train_loop" > gpurun_out/pytest_bn2.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_bn2.log
timeout 300 python tools/gpu_train_phases.py flat cls=ResNet loss=mse > gpurun_out/train_phases_resnet.log 2>&1; tail -14 gpurun_out/train_phases_resnet.log
