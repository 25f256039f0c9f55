#!/usr/bin/env python
"""A few launches of the fused residual-block kernel at the bench shape (for ncu)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from image_super_resolution_b200 import ops
from image_super_resolution_b200._lib import Context
dev = torch.device("cuda:0"); ctx = Context.get(0)
if os.environ.get("SRB_NO_PAIR") == "1":
    ctx.set_option(2, 1048576)          # SRB_PATH_NO_PAIR: single-CTA kernel
g = torch.Generator().manual_seed(0)
rb = ops.ResBlock(ctx, torch.randn(32, 64, 3, 3, generator=g) / 24, torch.zeros(32), torch.randn(64, 32, 3, 3, generator=g) / 17,
                  torch.zeros(64), "fp16")
N, H, W = 256, 64, 64
x = torch.randn(N, 8, H, W, 8, device=dev).half(); y = torch.empty_like(x)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 5):
    rb.forward(x, y, 0.2); rb.forward(y, x, 0.2)
torch.cuda.synchronize(); ctx.check_device()
