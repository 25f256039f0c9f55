#!/usr/bin/env python
"""Power / clocks while one kernel runs back to back for ~3 s (is the fused block power-limited?)."""
import os, sys, time, subprocess, threading
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from image_super_resolution_b200 import ops
from image_super_resolution_b200._lib import Context
dev = torch.device("cuda:0"); ctx = Context.get(0)
g = torch.Generator().manual_seed(0)
rb = ops.ResBlock(ctx, torch.randn(32, 64, 3, 3, generator=g) / 24, torch.zeros(32), torch.randn(64, 32, 3, 3, generator=g) / 17, torch.zeros(64), "fp16")
rc = ops.RowConv(ctx, torch.randn(64, 64, 3, 3, generator=g) / 24, torch.zeros(64), "fp16")
N, H, W = 256, 64, 64
x = torch.randn(N, 8, H, W, 8, device=dev).half(); y = torch.empty_like(x)
a = torch.randn(8192, 8192, device=dev, dtype=torch.bfloat16); b = torch.randn(8192, 8192, device=dev, dtype=torch.bfloat16)
def sample(stop, out):
    while not stop.is_set():
        r = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.sw_power_cap,clocks_event_reasons.hw_slowdown", "--format=csv,noheader"], capture_output=True, text=True)
        out.append(r.stdout.strip()); time.sleep(0.1)
def run(name, fn, flop, secs=3.0):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    stop, out = threading.Event(), []
    th = threading.Thread(target=sample, args=(stop, out)); th.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 0; t0 = time.time(); e0.record()
    while time.time() - t0 < secs:
        for _ in range(200): fn()
        n += 200
        torch.cuda.synchronize()
    e1.record(); torch.cuda.synchronize(); stop.set(); th.join()
    ms = e0.elapsed_time(e1) / n
    print(f"{name}: {ms * 1e3:.1f} us per launch, {flop / ms / 1e9:.0f} TFLOP/s; samples (clk, power, reasons, sw_power_cap, hw_slowdown):")
    for s in out[2:10]: print("   ", s)
P = N * H * W
run("fused residual block", lambda: rb.forward(x, y, 0.2), P * 73728.0)
run("row-fold conv1", lambda: rc.forward(x, y, x), P * 73728.0)
run("cuBLAS bf16 8192^3", lambda: torch.matmul(a, b), 2.0 * 8192 ** 3)
for skip, name in ((1, "no TMA"), (2, "no MMA"), (4, "no global I/O"), (6, "no MMA, no I/O"), (7, "protocol only")):
    ctx.debug_set(2, skip)
    for _ in range(3): rc.forward(x, y, x)
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): rc.forward(x, y, x)
    e1.record(); torch.cuda.synchronize()
    print(f"row-fold conv1, skip={skip} ({name}): {e0.elapsed_time(e1) / 20 * 1e3:.1f} us")
ctx.debug_set(2, 0)
