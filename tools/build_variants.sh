#!/bin/bash
# Build libsrb200.so variants with different fused-block ring sizes into build/v_<RH>_<RY>_<HSLOTS>_<XSTAGES>/ (sweeps on the GPU box).
set -e
cd "$(dirname "$0")/.."
for v in "$@"; do
  IFS=_ read rh ry hs xs <<< "$v"
  mkdir -p build/v_$v
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared \
    -DRB_RH=$rh -DRB_RY=$ry -DRB_HSLOTS=$hs -DRB_XSTAGES=$xs ${EXTRA_DEFS} \
    -o build/v_$v/libsrb200.so image_super_resolution_b200/csrc/srb200.cu &
done
wait
ls -la build/v_*/libsrb200.so
