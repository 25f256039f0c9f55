#!/bin/bash
# Round-1 probe logs (one gpurun call, no profiler): power/clock under sustained load, A/B of the kernel variants, tail / conv1
# bottleneck probes, fused-block probes, training phases.
set -u
mkdir -p gpurun_out
for t in gpu_power gpu_ab gpu_tail_probe gpu_rb_probe gpu_train_phases; do
    timeout 300 python tools/$t.py > gpurun_out/$t.log 2>&1; echo "$t: $?"
done
