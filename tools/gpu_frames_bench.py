#!/usr/bin/env python
"""BASELINE configs[2] on one GPU: 1080p frames through rs.upscale_frames into the headless byte sink (host frames in, BGR HWC
frames out, every copy inside the timed region).  Mpix/s = frames * 1080 * 1920 / time."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import image_super_resolution_b200 as srb
from image_super_resolution_b200 import rs
dev = torch.device("cuda:0")
model = srb.load(os.path.join(ROOT, "tests", "golden", "model_pt_weights.npz")).to(dev).eval()
g = torch.Generator().manual_seed(100)
for per_batch in (1, 2, 4):
    frames = [torch.randint(0, 256, (per_batch, 3, 1080, 1920), dtype=torch.uint8, generator=g).pin_memory() for _ in range(2)]
    batches = [frames[i % 2] for i in range(24 // per_batch)]
    rs.upscale_frames(model, batches[:3], rs.ByteSink())
    torch.cuda.synchronize(); t0 = time.perf_counter()
    sink = rs.ByteSink()
    n = rs.upscale_frames(model, batches, sink)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(f"{per_batch} frame(s) per batch: {n} frames in {dt * 1e3:7.1f} ms = {n * 1080 * 1920 / dt / 1e6:7.1f} Mpix/s ({n / dt:5.1f} fps), {sink.bytes} bytes written")
    # the unpipelined loop (one batch at a time: H2D, forward, blocking D2H) for comparison
    torch.cuda.synchronize(); t0 = time.perf_counter()
    sink2 = rs.ByteSink(); m = 0
    for b in batches:
        o = model.forward(b.to(dev, non_blocking=True), layout="hwc_bgr").cpu().numpy()
        for f in o:
            sink2.writeFrame(f); m += 1
    dt2 = time.perf_counter() - t0
    print(f"   unpipelined: {m * 1080 * 1920 / dt2 / 1e6:7.1f} Mpix/s")
