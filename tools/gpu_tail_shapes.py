#!/usr/bin/env python
"""Tail kernel: same pixel count in different shapes, full and load-only (DRAM locality of the C8-plane layout)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from image_super_resolution_b200 import ops
from image_super_resolution_b200._lib import Context
dev = torch.device("cuda:0"); ctx = Context.get(0)
g = torch.Generator().manual_seed(0)
t9 = ops.Tail9(ctx, torch.randn(3, 64, 9, 9, generator=g) / 72, torch.zeros(3), "fp16")
def timed(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
for N, H, W in [(256, 64, 64), (128, 64, 128), (64, 128, 128), (512, 64, 32), (1024, 32, 32), (16, 512, 128)]:
    x = torch.randn(N, 8, H, W, 8, device=dev).half()
    r = []
    for skip in (0, 2, 6):
        ctx.debug_set(2, skip); r.append(timed(lambda: t9.forward(x)))
    ctx.debug_set(2, 0)
    print(f"N={N:4d} H={H:3d} W={W:3d}: full {r[0]:6.1f} us, no MMA {r[1]:6.1f} us ({N*H*W*128/r[1]/1e6:.2f} TB/s), no MMA no stores {r[2]:6.1f} us")
