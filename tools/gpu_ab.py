#!/usr/bin/env python
"""A/B timing of model.pt forward at the bench shape with library debug switches (alternating, several rounds)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import image_super_resolution_b200 as srb
W = os.path.join(ROOT, "tests", "golden", "model_pt_weights.npz")
dev = torch.device("cuda:0")
model = srb.load(W).to(dev).eval()
x = torch.randint(0, 256, (256, 3, 64, 64), dtype=torch.uint8, device=dev); y = torch.empty_like(x)
ctx = model._ctx
variants = [(0, "current"), (32768, "head: PAIRDX + c8 pack"), (16384, "tail: tiled DXN"), (8192, "conv1: tiled"), (1024, "no programmatic dependent launch")]
if len(sys.argv) > 1: variants = [(int(a), "skip=" + a) for a in sys.argv[1:]]
def timed(n=20):
    for _ in range(3): model(x, out=y)
    torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): model(x, out=y)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
res = {k: [] for k, _ in variants}
for rnd in range(5):
    for k, _ in variants:
        ctx.debug_set(2, k); res[k].append(timed())
ctx.debug_set(2, 0)
for k, name in variants:
    v = sorted(res[k]); print(f"{name:36s}: median {v[len(v)//2]:7.1f} us  min {v[0]:7.1f}  max {v[-1]:7.1f}   -> {256*64*64/v[len(v)//2]:.0f} Mpix/s")
