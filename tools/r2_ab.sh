#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/gpu_ab_pair.py 2>&1 | tee gpurun_out/ab_pair.log
