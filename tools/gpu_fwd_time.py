#!/usr/bin/env python
"""model.pt forward at the bench shape: parity against the fp32 oracle on small batches, then device time of the whole
forward and of one fused residual block (CUDA events, 5 rounds of 20), CTA pairs against single CTAs (SRB_PATH_NO_PAIR)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import image_super_resolution_b200 as srb
from image_super_resolution_b200 import ops, _lib
from oracle import modelpt_oracle as mo
W = os.path.join(ROOT, "tests", "golden", "model_pt_weights.npz")
dev = torch.device("cuda:0")
model = srb.load(W).to(dev).eval()
ctx = model._ctx
g = torch.Generator().manual_seed(5)
for shape in ((2, 3, 40, 64), (4, 3, 24, 64), (1, 3, 33, 200), (3, 3, 17, 23), (5, 3, 9, 40), (1, 3, 3, 300)):
    xs = torch.randint(0, 256, shape, dtype=torch.uint8, generator=g)
    ref = mo.forward_u8(xs)
    for path, name in ((0, "pair"), (_lib.PATH_NO_PAIR, "single")):
        ctx.set_option(_lib.OPT_PATH, path)
        d = (model(xs.to(dev)).cpu().int() - ref.int()).abs()
        torch.cuda.synchronize()
        ok = int(d.max()) <= 2 and float((d == 0).float().mean()) >= 0.97
        print(f"parity {name:6s} {str(shape):18s} max {int(d.max()):3d} LSB exact {float((d == 0).float().mean()):.4f} {'ok' if ok else 'FAIL'}", flush=True)
    ctx.set_option(_lib.OPT_PATH, 0)
    try:
        ctx.check_device()
    except Exception as e:
        print("check_device:", e); sys.exit(1)
x = torch.randint(0, 256, (256, 3, 64, 64), dtype=torch.uint8, device=dev); y = torch.empty_like(x)
def timed(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
rb = ops.ResBlock(ctx, torch.randn(32, 64, 3, 3, generator=g) / 24, torch.zeros(32), torch.randn(64, 32, 3, 3, generator=g) / 17, torch.zeros(64), "fp16")
xa = torch.randn(256, 8, 64, 64, 8, device=dev).half(); ya = torch.empty_like(xa)
for rnd in range(2):
    for path, name in ((0, "pair"), (_lib.PATH_NO_PAIR, "single")):
        ctx.set_option(_lib.OPT_PATH, path)
        f = sorted(timed(lambda: model(x, out=y)) for _ in range(5))
        r = sorted(timed(lambda: rb.forward(xa, ya, 0.2)) for _ in range(5))
        print(f"{name:6s} forward median {f[2]:7.1f} us (min {f[0]:.1f}) = {256*64*64/f[2]:6.0f} Mpix/s   "
              f"resblock median {r[2]:6.1f} us (min {r[0]:.1f}) = {256*64*64*73728/r[2]/1e6:5.0f} TFLOP/s", flush=True)
ctx.set_option(_lib.OPT_PATH, 0)
ctx.check_device()
