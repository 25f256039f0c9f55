#!/bin/bash
# Build kernel-variant libraries (compile-time knobs) side by side: image_super_resolution_b200/variants/lib_<name>.so
# usage: tools/build_variants.sh name1:"-DRB_KSPLIT=0 -DRB_RH=4 ..." name2:"..."     (SRB200_LIB=... selects one at run time)
cd "$(dirname "$0")/../image_super_resolution_b200" || exit 1
mkdir -p variants
for spec in "$@"; do
    name="${spec%%:*}"; flags="${spec#*:}"
    ( nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 $flags -Xcompiler -fPIC -shared -o variants/lib_$name.so csrc/srb200.cu 2>&1 | grep -v "^$" | sed "s/^/[$name] /" ) &
done
wait
ls -la variants/
