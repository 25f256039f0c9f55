#!/bin/bash
mkdir -p gpurun_out
for cfg in "64 256" "63 260" "65 252" "64 256" "60 273"; do
  set -- $cfg
  echo "== H=$1 N=$2"
  SRB200_LIB=$PWD/image_super_resolution_b200/libsrb200_dbg.so timeout 300 python tools/gpu_rb_trace.py 0 $1 $2 > gpurun_out/rb_trace_h$1.log 2>&1
  grep -E "row period|total clk|SM clock|^launch 3|x stage free|x request|acc wait|ready -> issued|out full -> done" gpurun_out/rb_trace_h$1.log
done
