#!/bin/bash
V=$PWD/image_super_resolution_b200/variants
for r in 1 2; do for v in "$@"; do SRB200_LIB=$V/lib_$v.so timeout 200 python tools/gpu_side_kernels.py --quick 2>&1 | grep -E "QUICK|DIFFER|Error|error" | sed "s/^/$v /" | tee -a gpurun_out/side_ct_r2.log; done; done
