#!/usr/bin/env python
"""Sustained model.pt forward rate (the library named by SRB200_LIB): `secs` seconds of back-to-back forwards at the bench
shape after a 1 s heat-up, with NVML clock / power samples.  usage: python tools/gpu_sustained.py [secs]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import image_super_resolution_b200 as srb
sys.path.insert(0, ROOT)
from bench import ClockSampler
secs = float(sys.argv[1]) if len(sys.argv) > 1 else 3.0
dev = torch.device("cuda:0")
model = srb.load(os.path.join(ROOT, "tests", "golden", "model_pt_weights.npz")).to(dev).eval()
x = torch.randint(0, 256, (256, 3, 64, 64), dtype=torch.uint8, device=dev); y = torch.empty_like(x)
def run(seconds):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); n = 0
    e0.record()
    while time.perf_counter() - t0 < seconds:
        for _ in range(100): model(x, out=y)
        n += 100
    e1.record(); torch.cuda.synchronize()
    return n, e0.elapsed_time(e1)
run(1.0)
s = ClockSampler(0, period=0.02); s.start()
n, ms = run(secs)
c = s.stop()
print(f"{os.path.basename(os.environ.get('SRB200_LIB', 'libsrb200.so')):30s} sustained {n * 256 * 64 * 64 / ms / 1e3:7.1f} Mpix/s  ({ms / n * 1e3:6.1f} us/forward)  "
      f"SM {c['sm_mhz']} MHz  power median {c['power_w_median']:.0f} W  reasons {c['reasons']}", flush=True)
model._ctx.check_device()
