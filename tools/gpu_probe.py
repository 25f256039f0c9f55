#!/usr/bin/env python
"""Bottleneck probes for the conv kernel: time each layer type with parts of the pipeline disabled
(srb_debug_set key 2: 1 = no TMA tile loads, 2 = no MMAs, 4 = no global I/O in the epilogue)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from image_super_resolution_b200 import ops
from image_super_resolution_b200._lib import Context

dev = torch.device("cuda:0")
ctx = Context.get(0)
N, H, W = 256, 64, 64


def time_conv(cout, cin, k, residual, flags, iters=10):
    g = torch.Generator().manual_seed(0)
    w = torch.randn(cout, cin, k, k, generator=g) * 0.05
    b = torch.zeros(cout)
    pc = ops.PackedConv(ctx, w, b)
    pin = max(1, cin // 8)
    x = torch.randn((N, pin, H, W, 8), device=dev).half()
    if cout == 3:
        y = torch.zeros((N, 3, H, W), dtype=torch.uint8, device=dev)
        pout = 0
    else:
        pout = cout // 8
        y = torch.zeros((N, pout, H, W, 8), dtype=torch.float16, device=dev)
    res = torch.randn((N, pout, H, W, 8), device=dev).half() if residual else None
    out = {}
    for skip in (0, 1, 2, 4, 5, 6, 3, 7):
        ctx.debug_set(2, skip)
        for _ in range(2):
            ops.conv2d_fwd(pc, x, pin, 0, y, pout, 0, N, H, W, residual=res, res_planes=pout, flags=flags, slope=0.2)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            ops.conv2d_fwd(pc, x, pin, 0, y, pout, 0, N, H, W, residual=res, res_planes=pout, flags=flags, slope=0.2)
        e1.record()
        torch.cuda.synchronize()
        out[skip] = e0.elapsed_time(e1) / iters
    ctx.debug_set(2, 0)
    flop = 2.0 * cin * k * k * cout * N * H * W
    print(f"conv k{k} {cin}->{cout} res={residual}: " + "  ".join(
        f"skip{sk}={ms*1e3:.0f}us" for sk, ms in out.items()) + f"   full={flop/out[0]/1e9:.0f} TFLOP/s")


time_conv(32, 64, 3, False, ops.F_LRELU)
time_conv(64, 32, 3, True, 0)
time_conv(64, 64, 3, True, 0)
time_conv(64, 3, 9, False, ops.F_LRELU)
time_conv(3, 64, 9, False, ops.F_TANH | ops.F_U8OUT)
