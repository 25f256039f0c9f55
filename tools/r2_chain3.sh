#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/gpu_chain_trace.py 2 2>&1 | tee gpurun_out/chain_trace.log
