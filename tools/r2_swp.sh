#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/gpu_fwd_time.py 2>&1 | tee gpurun_out/fwd_time.log | tail -17
SRB200_LIB=$PWD/image_super_resolution_b200/libsrb200_dbg.so timeout 300 python tools/gpu_rb_trace.py > gpurun_out/rb_trace_swp.log 2>&1
grep -E "row period|total clk|SM clock|^launch 3|x stage free|x request|acc wait|ready -> issued" gpurun_out/rb_trace_swp.log
timeout 800 python tools/gpu_ab_variants.py --rounds 3 noswp base tail2 2>&1 | tee gpurun_out/ab_swp.log | tail -5
