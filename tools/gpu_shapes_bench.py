#!/usr/bin/env python
"""model.pt forward throughput at the shapes of BASELINE configs 1-3 (64x64 tiles, 1080p frames, a 4096^2 image), with the
row-fold kernels (default) and with the library's tiled layer-by-layer kernels (debug switch 8)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import image_super_resolution_b200 as srb

dev = torch.device("cuda:0")
model = srb.load(os.path.join(ROOT, "tests", "golden", "model_pt_weights.npz")).to(dev).eval()
ctx = model._ctx


def timed(x, y, reps):
    for _ in range(3):
        model.forward(x, out=y)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        model.forward(x, out=y)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for shape, reps in (((256, 3, 64, 64), 20), ((4, 3, 1080, 1920), 5), ((1, 3, 4096, 4096), 3), ((16, 3, 360, 640), 10), ((64, 3, 100, 100), 10)):
    x = torch.randint(0, 256, shape, dtype=torch.uint8, device=dev); y = torch.empty_like(x)
    px = shape[0] * shape[2] * shape[3]
    out = []
    for name, k in (("row-fold", 0), ("tiled", 8)):
        ctx.debug_set(2, k)
        ms = timed(x, y, reps)
        out.append(f"{name} {ms:8.3f} ms = {px / ms / 1e3:7.1f} Mpix/s")
    ctx.debug_set(2, 0)
    print(f"{str(shape):22s} " + "   ".join(out), flush=True)
    del x, y
    torch.cuda.empty_cache()
