#!/usr/bin/env python
"""tcgen05.mma micro-benchmark driver (see csrc/srb_ubench.cuh)."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from image_super_resolution_b200._lib import Context, lib, check
ctx = Context.get(0)
out = (C.c_longlong * 2)()
iters = 512
print("cycles per MMA per issuing warp (issue / complete); W = issuing warps, R = interleaved accumulators per warp")
for layout in (0,):
    for n in (16, 32, 64, 128, 256):
        for warps in (1, 2, 4):
            row = []
            for racc in (1, 2, 4, 8, 16):
                if racc * n * warps > 512:
                    continue
                check(lib().srb_ubench_mma(ctx.handle, n, layout, racc, iters, warps, 148, out))
                row.append(f"R{racc}: {out[0]/(iters*racc):6.1f}/{out[1]/(iters*racc):6.1f}")
            print(f"layout={layout} N={n:3d} W={warps}  " + "  ".join(row))
