#!/bin/bash
mkdir -p gpurun_out
SRB200_LIB=$PWD/image_super_resolution_b200/variants/lib_pair_noksplit.so timeout 300 python tools/gpu_fwd_time.py 2>&1 | tail -4 | tee gpurun_out/fwd_time_noksplit.log
SRB200_LIB=$PWD/image_super_resolution_b200/libsrb200_dbg.so timeout 300 python tools/gpu_rb_trace.py > gpurun_out/rb_trace_pair.log 2>&1; tail -34 gpurun_out/rb_trace_pair.log
