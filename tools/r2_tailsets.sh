#!/bin/bash
# tail9_kernel with 2 / 3 / 4 epilogue sets (variants from tools/build_variants.sh): parity tests of the tail, then cold-burst timings
V=$PWD/image_super_resolution_b200/variants
for v in t3 t4; do SRB200_LIB=$V/lib_$v.so timeout 300 python -m pytest tests/test_gpu_resblock.py -x -q -m gpu -k "tail" 2>&1 | tail -2; done
for r in 1 2; do for v in t2 t3 t4; do SRB200_LIB=$V/lib_$v.so timeout 200 python tools/gpu_side_kernels.py --quick 2>&1 | grep -E "QUICK|DIFFER|Error|error" | tee -a gpurun_out/side_tailsets_r2.log; done; done
