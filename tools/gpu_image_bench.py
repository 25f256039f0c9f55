#!/usr/bin/env python
"""BASELINE configs[3] on one GPU: one 4096x4096 uint8 image (host) through rs.upscale_image -> host result, per tile mode."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import image_super_resolution_b200 as srb
from image_super_resolution_b200 import rs
dev = torch.device("cuda:0")
model = srb.load(os.path.join(ROOT, "tests", "golden", "model_pt_weights.npz")).to(dev).eval()
g = torch.Generator().manual_seed(100)
img = torch.randint(0, 256, (3, 4096, 4096), dtype=torch.uint8, generator=g)
for mode, kw in (("whole", {}), ("halo", {}), ("halo", dict(tile=(1024, 2048))), ("halo", dict(tile=(2048, 2048))), ("reference", {})):
    rs.upscale_image(model, img[:, :512, :512].contiguous(), 96, mode, **kw)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    out = rs.upscale_image(model, img, 96, mode, **kw)
    dt = time.perf_counter() - t0
    buf = torch.empty((3, 4096, 4096), dtype=torch.uint8, pin_memory=True)
    rs.upscale_image(model, img, 96, mode, out=buf, **kw)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    out2 = rs.upscale_image(model, img, 96, mode, out=buf, **kw)
    dt2 = time.perf_counter() - t0
    assert torch.equal(out, out2)
    print(f"{mode:9s} {str(kw):28s} {dt * 1e3:8.1f} ms = {4096 * 4096 / dt / 1e6:7.1f} Mpix/s   into a reused pinned out: {dt2 * 1e3:6.1f} ms = {4096 * 4096 / dt2 / 1e6:7.1f} Mpix/s", flush=True)
