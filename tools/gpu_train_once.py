#!/usr/bin/env python
"""A few eager (no CUDA graph) forward + backward passes of the drop-in EResNet(16) at the C5 shape: the command the ncu
launch list of the training step is taken on (tools/ncu_summary.py aggregates it by kernel)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from image_super_resolution_b200 import models
from image_super_resolution_b200.training import L1Loss

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
B = 16
dev = torch.device("cuda:0")
torch.manual_seed(100)
net = models.EResNet(16, 0.2, 2).to(dev).train()
g = torch.Generator().manual_seed(100)
hr = (torch.rand(B, 3, 128, 128, generator=g) * 2 - 1).to(dev)
lr = torch.randn(B, 3, 64, 64, generator=g).to(dev)
loss_fn = L1Loss().to(dev)
for i in range(reps):
    net.zero_grad()
    with torch.autocast(device_type="cuda"):
        loss = loss_fn(net(lr), hr)
    if i == 0:
        net._train_plans[dev].use_graphs = False
    loss.backward()
torch.cuda.synchronize()
print("loss", loss.item())
