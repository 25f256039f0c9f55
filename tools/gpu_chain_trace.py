#!/usr/bin/env python
"""Timeline of CTA 0 of a conv chain (conv0 .. conv3 of an RDB as one launch) at the training shape, probe build only
(SRB200_LIB=.../libsrb200_dbg.so): clock64 stamps per pipeline step of the producer, the MMA issuer and the epilogue."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ.setdefault("SRB200_LIB", os.path.join(ROOT, "image_super_resolution_b200", "libsrb200_dbg.so"))
sys.path.insert(0, ROOT)
import torch
from image_super_resolution_b200 import ops
from image_super_resolution_b200._lib import Context, lib, check
dev = torch.device("cuda:0"); ctx = Context.get(0)
N, H, W = 16, 64, 64
ws = int(sys.argv[1]) if len(sys.argv) > 1 else 1
D = torch.randn(N, 24, H, W, 8, device=dev).to(torch.bfloat16)
pcs = []
for i in range(4):
    pk = ops.DeviceConv(ctx, 32, 64 + 32 * i, 3, "bf16", ws=ws)
    pk.fill(torch.randn(32, 64 + 32 * i, 3, 3, device=dev) * 0.02, torch.zeros(32, device=dev))
    pcs.append(pk)
chain = ops.ConvChain(ctx, [(pcs[i], ops.conv_args(D, 24, 0, D, 24, 8 + 4 * i, N, H, W, slope=0.01, flags=ops.F_LRELU)) for i in range(4)], keep=(D,))
def timed(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
def layerwise():
    for i in range(4):
        ops.conv2d_ex(pcs[i], D, 24, 0, D, 24, 8 + 4 * i, N, H, W, slope=0.01, flags=ops.F_LRELU)
print(f"ws={ws}: chain {timed(lambda: chain.run(dev)):6.1f} us   four launches {timed(layerwise):6.1f} us")
tr = torch.zeros(512, dtype=torch.int64, device=dev)
check(lib().srb_debug_trace(ctx.handle, tr.data_ptr())); chain.run(dev); torch.cuda.synchronize()
check(lib().srb_debug_trace(ctx.handle, None))
t8 = tr.cpu().view(64, 8); t0 = int(t8[0, 0])
print(f"CTA 0: entry->prologue {int(t8[1, 7]) - int(t8[0, 7])} clk, entry->exit {int(t8[2, 7]) - int(t8[0, 7])} clk")
print("step | tma_issue acc_free halo_ok mma_issued epi_wait acc_done epi_done | mma_len epi_len")
for i in range(16):
    if int(t8[i, 0]) == 0 and i > 0: break
    r = [int(v) - t0 for v in t8[i, :7]]
    print(f"{i:3d} | {r[0]:8d} {r[1]:8d} {r[2]:8d} {r[3]:8d} {r[4]:8d} {r[5]:8d} {r[6]:8d} | {r[3]-r[2]:6d} {r[6]-r[5]:6d}")
ctx.check_device()
