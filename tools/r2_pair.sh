#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/gpu_fwd_time.py 2>&1 | tee gpurun_out/fwd_time.log | tail -30
