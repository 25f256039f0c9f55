#!/usr/bin/env python
"""Timeline of CTA 0 of the fused residual-block kernel (clock64 stamps per row)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from image_super_resolution_b200 import ops
from image_super_resolution_b200._lib import Context, lib, check
dev = torch.device("cuda:0"); ctx = Context.get(0)
N, H, W = 256, 64, 64
g = torch.Generator().manual_seed(0)
rb = ops.ResBlock(ctx, torch.randn(32, 64, 3, 3, generator=g) / 24, torch.zeros(32), torch.randn(64, 32, 3, 3, generator=g) / 17,
                  torch.zeros(64), "fp16")
x = torch.randn(N, 8, H, W, 8, device=dev).half(); y = torch.empty_like(x)
skip = int(sys.argv[1]) if len(sys.argv) > 1 else 0
ctx.debug_set(2, skip)
tr = torch.zeros(1024 + 8 * 148, dtype=torch.int64, device=dev)
trs = [torch.zeros(1024 + 8 * 148, dtype=torch.int64, device=dev) for _ in range(4)]
for i in range(3):
    rb.forward(x, y, 0.2)
for i in range(4):                       # four traced launches back to back: launch-to-launch gaps from %globaltimer
    check(lib().srb_debug_trace(ctx.handle, trs[i].data_ptr()))
    rb.forward(x, y, 0.2)
check(lib().srb_debug_trace(ctx.handle, tr.data_ptr()))
rb.forward(x, y, 0.2)
torch.cuda.synchronize()
for i in range(4):
    c = trs[i].cpu()[1024:].view(148, 8)
    print(f"launch {i}: first CTA start {int(c[:,0].min()) - int(trs[0].cpu()[1024:].view(148,8)[:,0].min())} ns, last start +{int(c[:,0].max()-c[:,0].min())}, first end +{int(c[:,5].min()-c[:,0].min())}, last end +{int(c[:,5].max()-c[:,0].min())}")
check(lib().srb_debug_trace(ctx.handle, None)); ctx.debug_set(2, 0)
t = tr.cpu()[:1024].view(64, 16)
cta = tr.cpu()[1024:].view(148, 8)
t0 = int(t[0, 0])
print(f"skip={skip}: cycles relative to the first MMA iteration")
print("idx | mma: ready issued (len, period) | hid: wait full done (len) [ld, hsempty, sts, stwait, arrive] | out: wait full done (len) [ld, rfull, math, stwait, arrive]")
pm = None
for i in range(40):
    r = [int(v) - t0 if int(v) else 0 for v in t[i]]
    per = r[0] - pm if pm is not None else 0
    pm = r[0]
    print(f"{i:3d} | {r[0]:7d} {r[1]:7d} ({r[1]-r[0]:5d}, {per:5d}) | {r[2]:7d} {r[3]:7d} {r[4]:7d} ({r[4]-r[3]:5d}) [{r[11]-r[3]:4d} {r[12]-r[11]:4d} {r[13]-r[12]:4d} {r[14]-r[13]:4d} {r[4]-r[14]:4d}] | {r[5]:7d} {r[6]:7d} {r[7]:7d} ({r[7]-r[6]:5d}) [{r[10]-r[6]:4d} {r[8]-r[10]:4d} {r[15]-r[8]:4d} {r[9]-r[15]:4d} {r[7]-r[9]:4d}]")

g0 = int(cta[:, 0].min())
print("per CTA: launch skew ns | prologue clk | main clk (entry->all warps done) | producer done clk | end ns")
import statistics
pro = [int(c[2] - c[1]) for c in cta]; main = [int(c[4] - c[1]) for c in cta]; prod = [int(c[3] - c[1]) for c in cta]
st = [int(c[0]) - g0 for c in cta]; en = [int(c[5]) - g0 for c in cta]
for name, v in (("start ns", st), ("prologue clk", pro), ("producer done clk", prod), ("total clk", main), ("end ns", en)):
    print(f"  {name:18s} min {min(v):8d}  median {int(statistics.median(v)):8d}  max {max(v):8d}")
worst = sorted(range(148), key=lambda i: -main[i])[:6]
print("  slowest CTAs:", [(i, main[i]) for i in worst], " fastest:", sorted(main)[:4])
ghz = [float(c[4] - c[1]) / max(1.0, float(c[5] - c[0])) for c in cta]
print(f"  SM clock during the kernel (clock64 / globaltimer): min {min(ghz):.3f}  median {statistics.median(ghz):.3f}  max {max(ghz):.3f} GHz")
