#!/usr/bin/env python
"""Energy split of the fused residual block: sustained (2.5 s) time per launch, board power and SM clock for the probe variants of the
debug build (single-CTA kernel: SRB_PATH_NO_PAIR; skip bits of tools/gpu_rb_probe.py).  energy per launch = power x time.
usage: SRB200_LIB=.../libsrb200_dbg.so python tools/gpu_power_probe.py"""
import os, sys, time, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ.setdefault("SRB200_LIB", os.path.join(ROOT, "image_super_resolution_b200", "libsrb200_dbg.so"))
sys.path.insert(0, ROOT)
import torch
from image_super_resolution_b200 import ops, _lib
from image_super_resolution_b200._lib import Context
dev = torch.device("cuda:0"); ctx = Context.get(0)
g = torch.Generator().manual_seed(0)
rb = ops.ResBlock(ctx, torch.randn(32, 64, 3, 3, generator=g) / 24, torch.zeros(32), torch.randn(64, 32, 3, 3, generator=g) / 17, torch.zeros(64), "fp16")
N, H, W = 256, 64, 64
x = torch.randn(N, 8, H, W, 8, device=dev).half(); y = torch.empty_like(x)
def smi():
    r = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader,nounits"], capture_output=True, text=True)
    a, b = r.stdout.strip().split(",")
    return float(a), float(b)
time.sleep(2.0)
print("idle: %d MHz %.0f W" % smi())
fn = lambda: rb.forward(x, y, 0.2)
CASES = [(0, "full"), (2, "no MMA"), (16, "no convA MMAs"), (32, "no convB MMAs"), (1, "no TMA"), (4, "no global I/O in out-epilogue"),
         (6, "no MMA, no I/O"), (7, "protocol + epilogue ALU only"), (0, "full (again)")]
for pair in (False, True):
    ctx.set_option(_lib.OPT_PATH, 0 if pair else _lib.PATH_NO_PAIR)
    for skip, name in (CASES if not pair else [(0, "full, CTA pairs")]):
        ctx.debug_set(2, skip)
        for _ in range(3): fn()
        torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 0; t0 = time.time(); e0.record(); s = []
        while time.time() - t0 < 2.5:
            for _ in range(500): fn()
            n += 500
            if time.time() - t0 > 1.0: s.append(smi())
            torch.cuda.synchronize()
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / n * 1e3
        mhz = sorted(v[0] for v in s)[len(s) // 2]; w = sorted(v[1] for v in s)[len(s) // 2]
        print(f"skip={skip:3d} {name:34s} {us:7.1f} us/launch  {mhz:5.0f} MHz  {w:5.0f} W  -> {w * us * 1e-3:6.1f} mJ/launch  ({us * mhz * 1e-3:6.1f} kclk)", flush=True)
        time.sleep(1.0)
ctx.debug_set(2, 0); ctx.set_option(_lib.OPT_PATH, 0)
ctx.check_device()
