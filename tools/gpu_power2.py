#!/usr/bin/env python
"""Burst (20 launches) and sustained (2 s) time per launch of the fused block + power/clock, for wait-mode variants."""
import os, sys, time, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from image_super_resolution_b200 import ops
from image_super_resolution_b200._lib import Context
dev = torch.device("cuda:0"); ctx = Context.get(0)
g = torch.Generator().manual_seed(0)
rb = ops.ResBlock(ctx, torch.randn(32, 64, 3, 3, generator=g) / 24, torch.zeros(32), torch.randn(64, 32, 3, 3, generator=g) / 17, torch.zeros(64), "fp16")
rc = ops.RowConv(ctx, torch.randn(64, 64, 3, 3, generator=g) / 24, torch.zeros(64), "fp16")
N, H, W = 256, 64, 64
x = torch.randn(N, 8, H, W, 8, device=dev).half(); y = torch.empty_like(x)
def smi():
    r = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader"], capture_output=True, text=True)
    return r.stdout.strip()
for name, fn in (("fused block", lambda: rb.forward(x, y, 0.2)), ("row-fold conv1", lambda: rc.forward(x, y, x))):
    time.sleep(1.0)
    for _ in range(3): fn()
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): fn()
    e1.record(); torch.cuda.synchronize()
    burst = e0.elapsed_time(e1) / 20 * 1e3
    n = 0; t0 = time.time(); e0.record(); s = ""
    while time.time() - t0 < 2.0:
        for _ in range(500): fn()
        n += 500
        if not s and time.time() - t0 > 1.0: s = smi()
        torch.cuda.synchronize()
    e1.record(); torch.cuda.synchronize()
    print(f"  {name}: burst {burst:.1f} us, sustained {e0.elapsed_time(e1) / n * 1e3:.1f} us  [{s}]")
