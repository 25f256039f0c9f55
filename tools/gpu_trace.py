#!/usr/bin/env python
"""Per-step timeline of CTA 0 of one conv layer (clock64 stamps written by the kernel's three roles)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from image_super_resolution_b200 import ops
from image_super_resolution_b200._lib import Context, lib, check
dev = torch.device("cuda:0"); ctx = Context.get(0)
N, H, W = (int(sys.argv[1]) if len(sys.argv) > 1 else 256), 64, 64
GHZ = 1.9
CASES = ((32, 64, 3, 0), (64, 32, 3, 1), (64, 64, 3, 1), (64, 3, 9, 0), (3, 64, 9, 0))
if len(sys.argv) > 2 and sys.argv[2] == "dense":      # the dense blocks' weight-streaming convs
    CASES = ((32, 128, 3, 0), (32, 160, 3, 0), (64, 192, 3, 1))
for cout, cin, k, residual in CASES:
    g = torch.Generator().manual_seed(0)
    pc = ops.PackedConv(ctx, torch.randn(cout, cin, k, k, generator=g) * 0.05, torch.zeros(cout))
    pin = max(1, cin // 8)
    x = torch.randn((N, pin, H, W, 8), device=dev).half()
    if cout == 3:
        y = torch.zeros((N, 3, H, W), dtype=torch.uint8, device=dev); pout = 0; flags = ops.F_TANH | ops.F_U8OUT
    else:
        pout = cout // 8; y = torch.zeros((N, pout, H, W, 8), dtype=torch.float16, device=dev); flags = 0 if residual else ops.F_LRELU
    res = torch.randn((N, pout, H, W, 8), device=dev).half() if residual else None
    tr = torch.zeros(512, dtype=torch.int64, device=dev)
    for i in range(3):
        if i == 2:
            check(lib().srb_debug_trace(ctx.handle, tr.data_ptr()))
        ops.conv2d_fwd(pc, x, pin, 0, y, pout, 0, N, H, W, residual=res, res_planes=pout, flags=flags, slope=0.2)
    torch.cuda.synchronize()
    check(lib().srb_debug_trace(ctx.handle, None))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(20):
        ops.conv2d_fwd(pc, x, pin, 0, y, pout, 0, N, H, W, residual=res, res_planes=pout, flags=flags, slope=0.2)
    e1.record(); torch.cuda.synchronize()
    print(f"   back-to-back launch period {e0.elapsed_time(e1) / 20 * 1e3:.1f} us")
    t = tr.cpu().view(64, 8)
    t0 = int(t[0, 0])
    print(f"   CTA 0: entry -> prologue done {int(t[1, 7]) - int(t[0, 7])} clk, prologue done -> first TMA issue {t0 - int(t[1, 7])} clk, entry -> exit {int(t[2, 7]) - int(t[0, 7])} clk")
    print(f"== conv k{k} {cin}->{cout} res={residual}: stamps relative to first TMA issue (cycles)")
    print("step  tma_issue  acc_free  halo_ok  mma_issued  epi_wait  acc_done  epi_done | mma_issue_len  epi_len  step_period")
    prev = None
    for i in range(64):
        if int(t[i, 0]) == 0: break
        r = [int(v) - t0 for v in t[i, :7]]
        per = (r[6] - prev) if prev is not None else 0
        print(f"{i:3d} {r[0]:9d} {r[1]:9d} {r[2]:8d} {r[3]:10d} {r[4]:9d} {r[5]:9d} {r[6]:9d} | {r[3]-r[2]:8d} {r[6]-r[5]:8d} {per:8d}")
        prev = r[6]
