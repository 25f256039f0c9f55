#!/bin/bash
mkdir -p gpurun_out
python tools/gpu_rb_once.py 3 > gpurun_out/rb_once.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:resblock -s 4 -c 2 -o gpurun_out/ncu_rb_final -f python tools/gpu_rb_once.py 3 > gpurun_out/ncu_final.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/ncu_rb_final.ncu-rep
