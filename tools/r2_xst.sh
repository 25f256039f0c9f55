#!/bin/bash
# x-ring depth of the fused block: CTA pairs with 4 stages against single CTAs (72 KB of filters instead of 120 KB) with 5 / 6 stages
timeout 400 python tools/gpu_ab_variants.py --rounds 2 x4 2>&1 | tee -a gpurun_out/ab_xst_single_r2.log
timeout 600 python tools/gpu_ab_variants.py --rounds 2 --single s5 s6 2>&1 | tee -a gpurun_out/ab_xst_single_r2.log
