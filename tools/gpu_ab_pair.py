#!/usr/bin/env python
"""Interleaved A/B of the model.pt forward at the bench shape: fused residual block on CTA pairs (cta_group::2) against single
CTAs (SRB_PATH_NO_PAIR), alternating, 12 rounds of 40 forwards each, same box, same process."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import image_super_resolution_b200 as srb
from image_super_resolution_b200 import _lib
dev = torch.device("cuda:0")
model = srb.load(os.path.join(ROOT, "tests", "golden", "model_pt_weights.npz")).to(dev).eval()
ctx = model._ctx
x = torch.randint(0, 256, (256, 3, 64, 64), dtype=torch.uint8, device=dev); y = torch.empty_like(x)
def timed(n=40):
    for _ in range(3): model(x, out=y)
    torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): model(x, out=y)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
res = {0: [], _lib.PATH_NO_PAIR: []}
for rnd in range(12):
    for path in res:
        ctx.set_option(_lib.OPT_PATH, path); res[path].append(timed())
ctx.set_option(_lib.OPT_PATH, 0)
for path, name in ((0, "CTA pairs (cta_group::2)"), (_lib.PATH_NO_PAIR, "single CTAs")):
    v = sorted(res[path])
    print(f"{name:26s}: median {v[len(v)//2]:7.1f} us  min {v[0]:7.1f}  max {v[-1]:7.1f}  -> {256*64*64/v[len(v)//2]:.0f} Mpix/s   rounds: " + " ".join(f"{t:.0f}" for t in res[path]))
ctx.check_device()
