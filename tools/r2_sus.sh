#!/bin/bash
mkdir -p gpurun_out
for rnd in 1 2; do
for v in base park sleep200; do
  SRB200_LIB=$PWD/image_super_resolution_b200/variants/lib_$v.so timeout 120 python tools/gpu_sustained.py 3 2>&1 | tail -1
done; done | tee gpurun_out/sustained_variants.log
