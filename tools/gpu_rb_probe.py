#!/usr/bin/env python
"""Bottleneck probes of the fused residual-block kernel at the bench shape (256 x 64 x 64)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from image_super_resolution_b200 import ops
from image_super_resolution_b200._lib import Context
dev = torch.device("cuda:0"); ctx = Context.get(0)
N, H, W = 256, 64, 64
g = torch.Generator().manual_seed(0)
rb = ops.ResBlock(ctx, torch.randn(32, 64, 3, 3, generator=g) / 24, torch.zeros(32), torch.randn(64, 32, 3, 3, generator=g) / 17,
                  torch.zeros(64), "fp16")
x = torch.randn(N, 8, H, W, 8, device=dev).half()
y = torch.empty_like(x)
def timed(n=20):
    for _ in range(3): rb.forward(x, y, 0.2)
    torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): rb.forward(x, y, 0.2)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
for skip, name in [(0, "full"), (1024, "normal launches (no programmatic dependent launch)"), (1, "no TMA"), (2, "no MMA"), (4, "no global I/O in out-epilogue"), (3, "no TMA, no MMA"),
                   (64, "no TMEM re-zeroing (wrong results)"), (128, "no residual prefetch loads"), (256, "no output stores"), (130, "no MMA, no residual loads"), (258, "no MMA, no stores"), (512, "no MMA split at the ring wrap (wrong results)"), (16, "no convA MMAs"), (32, "no convB MMAs"), (6, "no MMA, no I/O"), (5, "no TMA, no I/O"), (7, "nothing but the barrier protocol + epilogue ALU")]:
    ctx.debug_set(2, skip)
    print(f"skip={skip} ({name}): {timed():7.1f} us")
ctx.debug_set(2, 0)
for grid in (148, 128, 74):
    ctx.debug_set(1, grid)
    print(f"grid={grid}: {timed():7.1f} us")
ctx.debug_set(1, 0)
try:
    ctx.check_device()
except Exception as e:
    print("check_device:", e)
