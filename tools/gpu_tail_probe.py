#!/usr/bin/env python
"""Bottleneck probes of the row-fold tail (9x9 64->3) and conv1 kernels at the bench shape."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from image_super_resolution_b200 import ops
from image_super_resolution_b200._lib import Context
dev = torch.device("cuda:0"); ctx = Context.get(0)
g = torch.Generator().manual_seed(0)
t9 = ops.Tail9(ctx, torch.randn(3, 64, 9, 9, generator=g) / 72, torch.zeros(3), "fp16")
rc = ops.RowConv(ctx, torch.randn(64, 64, 3, 3, generator=g) / 24, torch.zeros(64), "fp16")
def timed(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
for N in (256, 64):        # 64 tiles: 33.5 MB per tensor, every operand stays in the 126 MB L2 between calls
    H = W = 64
    x = torch.randn(N, 8, H, W, 8, device=dev).half(); y = torch.empty_like(x)
    for skip, name in [(0, "full"), (1, "no TMA"), (2, "no MMA"), (4, "no global stores"), (3, "no TMA, no MMA"), (6, "no MMA, no stores"), (7, "protocol + epilogue ALU only")]:
        ctx.debug_set(2, skip)
        print(f"N={N} skip={skip} ({name}): tail9 {timed(lambda: t9.forward(x)):6.1f} us   conv1 row-fold {timed(lambda: rc.forward(x, y, x)):6.1f} us", flush=True)
    ctx.debug_set(2, 0)
ctx.check_device()
