#!/usr/bin/env python
"""Fused residual-block kernel: time per launch over shapes with the same pixel count (DRAM locality probe)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from image_super_resolution_b200 import ops
from image_super_resolution_b200._lib import Context
dev = torch.device("cuda:0"); ctx = Context.get(0)
g = torch.Generator().manual_seed(0)
rb = ops.ResBlock(ctx, torch.randn(32, 64, 3, 3, generator=g) / 24, torch.zeros(32), torch.randn(64, 32, 3, 3, generator=g) / 17,
                  torch.zeros(64), "fp16")
def timed(x, y, n=20):
    for _ in range(3): rb.forward(x, y, 0.2)
    torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): rb.forward(x, y, 0.2)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
for N, H, W in [(256, 64, 64), (128, 64, 128), (64, 128, 128), (512, 64, 32), (128, 128, 64), (1024, 64, 16), (32, 256, 128), (1, 64, 64), (2, 64, 64), (4, 64, 64), (8, 64, 64), (16, 64, 64), (18, 64, 64), (19, 64, 64), (37, 64, 64), (74, 64, 64), (148, 64, 64)]:
    x = torch.randn(N, 8, H, W, 8, device=dev).half(); y = torch.empty_like(x)
    t = timed(x, y)
    px = N * H * W
    print(f"N={N:4d} H={H:3d} W={W:3d}: {t:7.1f} us  {px / t:8.1f} px/us  {px * 73728 / t / 1e6:6.1f} TFLOP/s  HBM {px * 256 / t / 1e6:5.2f} TB/s")
ctx.check_device()
