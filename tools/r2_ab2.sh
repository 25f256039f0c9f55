#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python tools/gpu_ab_variants.py --rounds 2 base late nogm sleep0 sleep100 2>&1 | tee gpurun_out/ab_cold_variants.log
timeout 300 python tools/gpu_ab_variants.py --rounds 1 --single base 2>&1 | tee -a gpurun_out/ab_cold_variants.log
