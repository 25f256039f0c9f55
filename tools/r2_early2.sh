#!/bin/bash
mkdir -p gpurun_out
for rnd in 1 2; do for v in late early; do
  echo "== $v (round $rnd)"
  SRB200_LIB=$PWD/image_super_resolution_b200/variants/lib_${v}_dbg.so timeout 300 python tools/gpu_rb_trace.py > gpurun_out/rb_trace_${v}_$rnd.log 2>&1
  grep -E "row period|total clk|SM clock|^launch 3|out full -> done|hidden full -> published|convA ready -> issued|acc wait" gpurun_out/rb_trace_${v}_$rnd.log
done; done
