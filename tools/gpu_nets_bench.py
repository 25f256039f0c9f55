#!/usr/bin/env python
"""Inference throughput of the class nets (EResNet(16) x2, utils/models.py:621-650) through the drop-in classes: eager
launches vs the captured forward (models._NetPlan.use_graphs).  Output Mpix/s = N * (2H) * (2W) / time."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from image_super_resolution_b200 import models
dev = torch.device("cuda:0")
torch.manual_seed(100)
net = models.EResNet(16, 0.2, 2).to(dev).eval()
def timed(x, n=10):
    with torch.no_grad():
        for _ in range(3): net(x)
        torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(n): net(x)
        b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
for shape in ((1, 3, 96, 96), (16, 3, 64, 64), (64, 3, 64, 64), (1, 3, 540, 960)):
    x = torch.randn(shape, device=dev)
    res = []
    for g in (False, True):
        net._plan(dev).use_graphs = g
        ms = timed(x)
        res.append(f"{'graph' if g else 'eager'} {ms:8.3f} ms = {shape[0] * shape[2] * shape[3] * 4 / ms / 1e3:7.1f} Mpix/s out")
    a = net(x); net._plan(dev).use_graphs = False; b = net(x)
    print(f"{str(shape):18s} " + "   ".join(res) + f"   max|graph - eager| = {(a - b).abs().max().item():.1e}", flush=True)
