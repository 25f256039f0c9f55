#!/usr/bin/env python
"""BASELINE configs[2] and [3] from ONE host process over G = 1, 2, 4, 8 devices (the case where the host is the limiter):
  frames: 64 x 1080p uint8 frames (pinned host) through rs.upscale_frames(ShardedModel) into the byte sink, BGR HWC out;
  image:  one 4096 x 4096 image through rs.upscale_image(ShardedModel), 2 x 4 grid of 2048 x 1024 halo tiles.
Results are checked bit-equal to G = 1.  usage: python tools/gpu_scale_rs.py [max_devices]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import image_super_resolution_b200 as srb
from image_super_resolution_b200 import rs
W = os.path.join(ROOT, "tests", "golden", "model_pt_weights.npz")
ndev = min(int(sys.argv[1]) if len(sys.argv) > 1 else 8, torch.cuda.device_count())
Gs = [g for g in (1, 2, 4, 8) if g <= ndev]
g = torch.Generator().manual_seed(100)
PER = 2
frames = [torch.randint(0, 256, (PER, 3, 1080, 1920), dtype=torch.uint8, generator=g).pin_memory() for _ in range(4)]
img = torch.randint(0, 256, (3, 4096, 4096), dtype=torch.uint8, generator=g).pin_memory()
out = torch.empty((3, 4096, 4096), dtype=torch.uint8, pin_memory=True)
base = srb.load(W)
ref_frames = ref_img = None
f1 = i1 = None
for G in Gs:
    sm = rs.ShardedModel(base, devices=list(range(G)))
    nb = 32 * max(1, G // 2)                                   # 64 frames at G <= 2, more batches for more devices (weak scaling of the run length)
    batches = [frames[i % 4] for i in range(nb)]
    rs.upscale_frames(sm, batches[:6 * G], rs.ByteSink(copy=False))        # warm-up: every slot of every replica's ring captured
    rs.upscale_frames(sm, batches[:6 * G], rs.ByteSink(copy=False), threads=not (G > 1))
    for copy, thr in ((False, True), (False, False), (True, True), (True, False)):
        sink = rs.ByteSink(copy=copy)
        for d in range(G): torch.cuda.synchronize(d)
        t0 = time.perf_counter()
        n = rs.upscale_frames(sm, batches, sink, threads=thr)
        for d in range(G): torch.cuda.synchronize(d)
        dt = time.perf_counter() - t0
        v = n * 1080 * 1920 / dt / 1e6
        if not copy and thr:
            f1 = f1 or v
        print(f"frames G={G} {'feeder thread per device' if thr else 'single host thread      '} sink {'tobytes (writer stand-in)' if copy else 'count-only              '}: "
              f"{n} frames {dt * 1e3:8.1f} ms = {v:8.1f} Mpix/s = {n / dt:7.1f} fps" + (f"  scaling {v / f1:.2f}x" if (not copy and thr) else ""), flush=True)
    keep = rs.ByteSink(keep=True)
    rs.upscale_frames(sm, batches[:2 * G + 1], keep)
    got = np.stack(keep.kept)
    if ref_frames is None:
        ref_frames = got[:6]
    assert np.array_equal(got[:6], ref_frames), "frames differ from G = 1"
    for rep in range(3):
        t0 = time.perf_counter()
        o = rs.upscale_image(sm, img, 96, "halo", tile=(2048, 1024), out=out)
        dt = time.perf_counter() - t0
    i1 = i1 or dt
    if ref_img is None:
        ref_img = o.clone()
    assert torch.equal(o, ref_img), "image differs from G = 1"
    for rep in range(3):
        t0 = time.perf_counter(); o3 = rs.upscale_image(sm, img, 96, "halo", tile=(1024, 1024), out=out); dt3 = time.perf_counter() - t0
    assert torch.equal(o3, ref_img)
    t0 = time.perf_counter(); o2 = rs.upscale_image(sm, img, 96, "reference", out=out); dt2 = time.perf_counter() - t0
    t0 = time.perf_counter(); o2 = rs.upscale_image(sm, img, 96, "reference", out=out); dt2 = time.perf_counter() - t0
    print(f"image  G={G} 4096^2, 2x4 halo tiles: {dt * 1e3:7.2f} ms = {4096 * 4096 / dt / 1e6:8.1f} Mpix/s  scaling {i1 / dt:.2f}x   (in 48 MB + out 48 MB over PCIe)"
          f"   4x4 tiles of 1024^2: {dt3 * 1e3:6.2f} ms   reference windows (1849 x 96^2): {dt2 * 1e3:7.2f} ms", flush=True)
    del sm
print("results bit-equal to G = 1: frames yes, image yes")
