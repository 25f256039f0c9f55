#!/bin/bash
mkdir -p gpurun_out
python tools/gpu_rb_once.py 3 > gpurun_out/rb_once.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:resblock -s 4 -c 2 -o gpurun_out/ncu_rb_pair -f python tools/gpu_rb_once.py 3 > gpurun_out/ncu_pair.log 2>&1
echo "ncu pair rc=$?"
SRB_NO_PAIR=1 ncu --set full --clock-control none --import-source on -k regex:resblock -s 4 -c 2 -o gpurun_out/ncu_rb_single -f python tools/gpu_rb_once.py 3 > gpurun_out/ncu_single.log 2>&1
echo "ncu single rc=$?"
ls -la gpurun_out/*.ncu-rep
