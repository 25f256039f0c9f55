#!/bin/bash
mkdir -p gpurun_out
timeout 800 python tools/gpu_ab_variants.py --rounds 3 late early 2>&1 | tee gpurun_out/ab_early.log
SRB200_LIB=$PWD/image_super_resolution_b200/libsrb200_dbg.so timeout 300 python tools/gpu_rb_trace.py > gpurun_out/rb_trace_early.log 2>&1; tail -32 gpurun_out/rb_trace_early.log
timeout 300 python tools/gpu_fwd_time.py 2>&1 | tee gpurun_out/fwd_time.log | tail -17
