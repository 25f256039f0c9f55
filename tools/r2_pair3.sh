#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/gpu_fwd_time.py 2>&1 | tee gpurun_out/fwd_time.log | tail -17
SRB200_LIB=$PWD/image_super_resolution_b200/libsrb200_dbg.so timeout 300 python tools/gpu_rb_trace.py > gpurun_out/rb_trace_pair.log 2>&1; tail -30 gpurun_out/rb_trace_pair.log | head -18
