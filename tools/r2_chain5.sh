#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_train.py tests/test_gpu_optim.py -m gpu -x -q > gpurun_out/pytest_train.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_train.log | cut -c1-300
for cfg in "train_chain5=1" "train_chain5=0" "train_chain5=1" "train_chain5=0"; do
  echo "== $cfg"; timeout 300 python tools/gpu_train_phases.py flat $cfg 2>&1 | grep -E "forward|backward|un-sync"
done | tee gpurun_out/chain5.log
