#!/usr/bin/env python
"""The batched weight-gradient launch of one EResNet(16) training step (C5 shape): time per split-K factor and with
parts of the pipeline disabled (srb_debug_set key 2: 1 = no TMA loads, 2 = no MMAs)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from image_super_resolution_b200 import models
from image_super_resolution_b200.training import L1Loss
from image_super_resolution_b200._lib import Context
B = 16
dev = torch.device("cuda:0"); ctx = Context.get(0)
torch.manual_seed(100)
net = models.EResNet(16, 0.2, 2).to(dev).train()
g = torch.Generator().manual_seed(100)
hr = (torch.rand(B, 3, 128, 128, generator=g) * 2 - 1).to(dev)
lr = torch.randn(B, 3, 64, 64, generator=g).to(dev)
loss_fn = L1Loss().to(dev)
with torch.autocast(device_type="cuda"):
    loss = loss_fn(net(lr), hr)
loss.backward()
plan = net._train_plans[dev]
gb = plan.grad_batch
print("units", gb.units)
def timed(fn, n=10):
    for _ in range(2): fn()
    torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
for skip in (0, 1, 2, 3):
    ctx.debug_set(2, skip)
    print(f"skip={skip}:", "  ".join(f"splitk {k}: {timed(lambda: gb.run(dev, k)):7.1f} us" for k in (1, 2, 3, 4, 8)))
ctx.debug_set(2, 131072)            # units in job order (RDB-major: siblings share their X / dY planes through L2)
from image_super_resolution_b200 import ops
gb2 = ops.GradBatch(ctx, gb._keep, "bf16")
ctx.debug_set(2, 0)
print("job order:", "  ".join(f"splitk {k}: {timed(lambda: gb2.run(dev, k)):7.1f} us" for k in (1, 2, 3)))
