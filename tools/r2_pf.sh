#!/bin/bash
mkdir -p gpurun_out
timeout 800 python tools/gpu_ab_variants.py --rounds 3 base pf2 pf4 pf8 2>&1 | tee gpurun_out/ab_pf.log
SRB200_LIB=image_super_resolution_b200/libsrb200_dbg.so timeout 300 python tools/gpu_ubench_mix.py --pair > gpurun_out/ubench_mix_r2.log 2>&1; echo "ubench rc=$?"; tail -32 gpurun_out/ubench_mix_r2.log
