#!/usr/bin/env python
"""tcgen05.mma micro-benchmark of the fused-block operand layouts (see csrc/srb_ubench.cuh)."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from image_super_resolution_b200._lib import Context, lib, check
ctx = Context.get(0)
out = (C.c_longlong * 2)()
iters = 512
print("cycles per MMA (issue / complete); layout 0 = conv halo tile, 3 = x-row boxes (LBO 160, SBO 1280), 4 = hidden-row boxes (SBO 640), 5 = 11-px boxes")
for n in (32, 64, 96, 128, 192):
    for layout in (0, 3, 4, 5):
        row = []
        for racc in (1, 2, 4):
            if racc * n > 512:
                continue
            check(lib().srb_ubench_mma(ctx.handle, n, layout, racc, iters, 1, 148, out))
            row.append(f"R{racc}: {out[0]/(iters*racc):6.1f}/{out[1]/(iters*racc):6.1f}")
        print(f"N={n:3d} layout={layout}  " + "  ".join(row))
