#!/bin/bash
# probes of the tail / conv1 kernels (debug library) and an A/B of deeper operand rings (variants built by tools/build_variants.sh)
V=$PWD/image_super_resolution_b200/variants
SRB200_LIB=$PWD/image_super_resolution_b200/libsrb200_dbg.so timeout 300 python tools/gpu_tail_probe.py 2>&1 | tee gpurun_out/tail_probe_r2.log
for r in 1 2; do for v in base2 st; do SRB200_LIB=$V/lib_$v.so timeout 200 python tools/gpu_side_kernels.py --quick 2>&1 | grep -E "QUICK|DIFFER|Error|error" | tee -a gpurun_out/side_stages_r2.log; done; done
