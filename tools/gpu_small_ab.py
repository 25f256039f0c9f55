#!/usr/bin/env python
"""A/B of the small-problem tiling (two-tile super-tiles) on the training step: forward + backward time of EResNet(16), B = 16,
with debug bits 0 (small tiling everywhere), 262144 (large tiling for the weight-streaming kernels), 65536 (large everywhere)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from image_super_resolution_b200 import models
from image_super_resolution_b200.training import L1Loss
from image_super_resolution_b200._lib import Context
B = 16
dev = torch.device("cuda:0"); ctx = Context.get(0)
g = torch.Generator().manual_seed(100)
hr = (torch.rand(B, 3, 128, 128, generator=g) * 2 - 1).to(dev)
lr = torch.randn(B, 3, 64, 64, generator=g).to(dev)
loss_fn = L1Loss().to(dev)
for bits in (0, 262144, 65536, 0, 262144):
    ctx.debug_set(2, bits)
    torch.manual_seed(100)
    net = models.EResNet(16, 0.2, 2).to(dev).train()
    def step():
        net.zero_grad()
        with torch.autocast(device_type="cuda"):
            loss = loss_fn(net(lr), hr)
        loss.backward()
    for _ in range(3): step()                     # captures the graphs with this tiling
    torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10): step()
    b.record(); torch.cuda.synchronize()
    print(f"debug bits {bits:6d}: fwd+bwd {a.elapsed_time(b) / 10:.3f} ms")
ctx.debug_set(2, 0)
