#!/usr/bin/env python
"""Per-kernel timing of the backward building blocks at the C5 shape (B=16, 64x64 LR)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from image_super_resolution_b200 import ops
from image_super_resolution_b200._lib import Context
dev = torch.device("cuda:0"); ctx = Context.get(0)
N, H, W = 16, 64, 64
def timed(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
D = torch.randn(N, 24, H, W, 8, device=dev).to(torch.bfloat16)
G = torch.randn(N, 24, H, W, 8, device=dev).to(torch.bfloat16)
for cin, cout in [(64, 32), (96, 32), (128, 32), (160, 32), (192, 64)]:
    dw = torch.zeros(cout, cin, 3, 3, device=dev)
    t = timed(lambda: ops.conv2d_wgrad(D, 24, 0, cin, G, 24, 8, cout, dw, N, H, W))
    fl = 2 * N * H * W * cin * cout * 9
    print(f"wgrad {cin:3d}->{cout:2d}: {t:7.1f} us  {fl / t / 1e6:6.1f} TFLOP/s")
db = torch.zeros(64, device=dev)
print(f"bias_grad 32: {timed(lambda: ops.bias_grad(G, 24, 8, 32, db, N, H, W)):7.1f} us")
print(f"bias_grad 64: {timed(lambda: ops.bias_grad(G, 24, 0, 64, db, N, H, W)):7.1f} us")
w = torch.randn(32, 192, 3, 3, device=dev) * 0.02
for cin_, cout_ in [(32, 64), (32, 32), (64, 64)]:
    pk = ops.DeviceConv(ctx, cout_, cin_, 3, "bf16")
    pk.fill(torch.randn(cin_, cout_, 3, 3, device=dev) * 0.02, transpose=True)
    t = timed(lambda: ops.conv2d_ex(pk, G, 24, 8, G, 24, 0, N, H, W, residual=G, res_planes=24, res_off=0))
    fl = 2 * N * H * W * cin_ * cout_ * 9
    print(f"dgrad {cin_}->{cout_} in-place acc: {t:7.1f} us  {fl / t / 1e6:6.1f} TFLOP/s")
    t = timed(lambda: ops.conv2d_ex(pk, G, 24, 8, G, 24, 0, N, H, W, residual=G, res_planes=24, res_off=0, mask=D, mask_planes=24,
                                    mask_off=0, mask_p0=0, mask_p1=cout_ // 8, mask_slope=0.01))
    print(f"   + mask: {t:7.1f} us")
for cin_ in (64, 96, 128, 160, 192):
    pk = ops.DeviceConv(ctx, 64 if cin_ == 192 else 32, cin_, 3, "bf16")
    pk.fill(torch.randn(pk.cout, cin_, 3, 3, device=dev) * 0.02)
    t = timed(lambda: ops.conv2d_ex(pk, D, 24, 0, G, 24, 16, N, H, W, slope=0.01, flags=ops.F_LRELU))
    fl = 2 * N * H * W * cin_ * pk.cout * 9
    print(f"fwd {cin_}->{pk.cout}: {t:7.1f} us  {fl / t / 1e6:6.1f} TFLOP/s")
pk = ops.DeviceConv(ctx, 32, 64, 3, "bf16")
wsrc = torch.randn(32, 64, 3, 3, device=dev)
print(f"pack (eager, host-bound): {timed(lambda: pk.fill(wsrc)):7.1f} us")
x9 = torch.randn(N, 1, 2 * H, 2 * W, 8, device=dev).to(torch.bfloat16)
u9 = torch.randn(N, 8, 2 * H, 2 * W, 8, device=dev).to(torch.bfloat16)
dw9 = torch.zeros(3, 64, 9, 9, device=dev)
print(f"wgrad9 tail (HR): {timed(lambda: ops.conv9_wgrad(u9, 8, 0, 64, x9, 1, 0, 3, dw9, N, 2 * H, 2 * W)):7.1f} us")
dw9h = torch.zeros(64, 3, 9, 9, device=dev)
print(f"wgrad9 head (LR): {timed(lambda: ops.conv9_wgrad(D, 24, 23, 3, G, 24, 0, 64, dw9h, N, H, W)):7.1f} us")
Context.get(0).check_device()
