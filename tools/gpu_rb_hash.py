#!/usr/bin/env python
"""Hashes of the fused residual block's output on several shapes (one line per shape): run with two libraries (SRB200_LIB) and
diff the output to check that a kernel variant is bit-identical."""
import os, sys, hashlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from image_super_resolution_b200 import ops
from image_super_resolution_b200._lib import Context
dev = torch.device("cuda:0"); ctx = Context.get(0)
g = torch.Generator().manual_seed(0)
rb = ops.ResBlock(ctx, torch.randn(32, 64, 3, 3, generator=g) / 24, torch.randn(32, generator=g) * 0.1,
                  torch.randn(64, 32, 3, 3, generator=g) / 17, torch.randn(64, generator=g) * 0.1, "fp16")
for (N, H, W) in ((256, 64, 64), (6, 64, 64), (5, 40, 56), (3, 33, 40), (2, 24, 200), (1, 50, 136), (2, 1080, 1920), (7, 9, 88), (4, 3, 64), (2, 2, 64)):
    x = (torch.randn(N, 8, H, W, 8, generator=g) * 0.5).half().to(dev); y = torch.empty_like(x)
    rb.forward(x, y, 0.2); torch.cuda.synchronize()
    print(N, H, W, hashlib.sha1(y.cpu().numpy().tobytes()).hexdigest())
ctx.check_device()
