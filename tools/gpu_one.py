#!/usr/bin/env python
"""Run one conv layer type a few times at config-2 size (for ncu captures)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from image_super_resolution_b200 import ops
from image_super_resolution_b200._lib import Context
cout, cin, k, residual = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
iters = int(sys.argv[5]) if len(sys.argv) > 5 else 3
dev = torch.device("cuda:0"); ctx = Context.get(0)
N, H, W = 256, 64, 64
g = torch.Generator().manual_seed(0)
pc = ops.PackedConv(ctx, torch.randn(cout, cin, k, k, generator=g) * 0.05, torch.zeros(cout))
pin = max(1, cin // 8)
x = torch.randn((N, pin, H, W, 8), device=dev).half()
if cout == 3:
    y = torch.zeros((N, 3, H, W), dtype=torch.uint8, device=dev); pout = 0; flags = ops.F_TANH | ops.F_U8OUT
else:
    pout = cout // 8; y = torch.zeros((N, pout, H, W, 8), dtype=torch.float16, device=dev); flags = 0 if residual else ops.F_LRELU
res = torch.randn((N, pout, H, W, 8), device=dev).half() if residual else None
ctx.debug_set(2, int(os.environ.get('SRB_SKIP', '0')))
N = int(os.environ.get('SRB_N', N))
for _ in range(iters):
    ops.conv2d_fwd(pc, x, pin, 0, y, pout, 0, N, H, W, residual=res, res_planes=pout, flags=flags, slope=0.2)
ctx.check_device()
print("ok")
