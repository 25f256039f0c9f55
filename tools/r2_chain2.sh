#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_train.py -m gpu -x -q 2>&1 | tail -2
for cfg in "train_conv_fold=1" "train_conv_fold=2" "train_conv_fold=1"; do
  echo "== $cfg"; timeout 300 python tools/gpu_train_phases.py flat $cfg 2>&1 | grep -E "forward|backward|un-sync"
done | tee gpurun_out/chain_variants.log
