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
if len(sys.argv) > 2: H = int(sys.argv[2])
if len(sys.argv) > 3: N = int(sys.argv[3])
g = torch.Generator().manual_seed(0)
rb = ops.ResBlock(ctx, torch.randn(32, 64, 3, 3, generator=g) / 24, torch.zeros(32), torch.randn(64, 32, 3, 3, generator=g) / 17,
                  torch.zeros(64), "fp16")
x = torch.randn(N, 8, H, W, 8, device=dev).half(); y = torch.empty_like(x)
skip = int(sys.argv[1]) if len(sys.argv) > 1 else 0
ctx.debug_set(2, skip)
tr = torch.zeros(4096, dtype=torch.int64, device=dev)      # [32 rows][32 stamps] | per-CTA [148][8] at 1024 | per-launch records at 2300
trs = [torch.zeros(4096, dtype=torch.int64, device=dev) for _ in range(4)]
for i in range(3):
    rb.forward(x, y, 0.2)
for i in range(4):                       # four traced launches back to back: launch-to-launch gaps from %globaltimer
    check(lib().srb_debug_trace(ctx.handle, trs[i].data_ptr()))
    rb.forward(x, y, 0.2)
check(lib().srb_debug_trace(ctx.handle, tr.data_ptr()))
rb.forward(x, y, 0.2)
torch.cuda.synchronize()
for i in range(4):
    c = trs[i].cpu()[1024:1024 + 8 * 148].view(148, 8)
    print(f"launch {i}: first CTA start {int(c[:,0].min()) - int(trs[0].cpu()[1024:1024 + 8 * 148].view(148,8)[:,0].min())} ns, last start +{int(c[:,0].max()-c[:,0].min())}, first end +{int(c[:,5].min()-c[:,0].min())}, last end +{int(c[:,5].max()-c[:,0].min())}")
check(lib().srb_debug_trace(ctx.handle, None)); ctx.debug_set(2, 0)
t = tr.cpu()[:1024].view(32, 32)
cta = tr.cpu()[1024:1024 + 8 * 148].view(148, 8)
t0 = int(t[0, 19])
print(f"skip={skip}: CTA 0, cycles relative to the first x-row request; row i = x row i = hidden row i = output row i")
print("row | x: free req'd arrived | convA: ready issued (period) | hid: full done [ld+zero, hsempty, sts, stwait, arrive] |"
      " convB: hsfull acq issued | out: wait full done (len) [ld+zero, ldg, math+stg, stwait, arrive]")
pm = None
for i in range(32):
    r = [int(v) - t0 if int(v) else 0 for v in t[i]]
    per = r[0] - pm if pm is not None else 0
    pm = r[0]
    print(f"{i:3d} | {r[19]:6d} {r[20]:6d} {r[21]:6d} | {r[0]:6d} {r[1]:6d} ({per:5d}) | {r[3]:6d} {r[4]:6d} [{r[11]-r[3]:4d} {r[12]-r[11]:4d} {r[13]-r[12]:4d} {r[14]-r[13]:4d} {r[4]-r[14]:4d}] |"
          f" {r[16]:6d} {r[17]:6d} {r[18]:6d} | {r[5]:6d} {r[6]:6d} {r[7]:6d} ({r[7]-r[6]:5d}) [{r[10]-r[6]:4d} {r[8]-r[10]:4d} {r[15]-r[8]:4d} {r[9]-r[15]:4d} {r[7]-r[9]:4d}]")
# steady-state summary over rows 8..30
import statistics as _st
def med(f):
    return int(_st.median([f([int(v) for v in t[i]], [int(v) for v in t[i + 1]]) for i in range(8, 30)]))
print("steady state (median over rows 8..29, clk):")
print(f"  row period (convA ready to ready)        {med(lambda a, b: b[0] - a[0])}")
print(f"  x request -> arrival                     {med(lambda a, b: a[21] - a[20])}")
print(f"  x arrival -> convA ready (acc wait)      {med(lambda a, b: a[0] - a[21])}")
print(f"  convA ready -> issued                    {med(lambda a, b: a[1] - a[0])}")
print(f"  convA(r+1) issued -> hidden row r full   {med(lambda a, b: a[3] - b[1])}")
print(f"  hidden full -> published                 {med(lambda a, b: a[4] - a[3])}")
print(f"  hidden published -> convB sees it        {med(lambda a, b: a[16] - a[4])}")
print(f"  convB sees hidden -> acc acquired        {med(lambda a, b: a[17] - a[16])}")
print(f"  convB acquired -> issued                 {med(lambda a, b: a[18] - a[17])}")
print(f"  convB(h+1) issued -> out row h full      {med(lambda a, b: a[6] - b[18])}")
print(f"  out full -> done                         {med(lambda a, b: a[7] - a[6])}")
print(f"  out-epilogue idle (done -> next full)    {med(lambda a, b: b[6] - a[7])}")
print(f"  x stage free -> requested                {med(lambda a, b: a[20] - a[19])}")
print(f"  x row r requested at row period offset   {med(lambda a, b: a[0] - a[20])} clk before convA(r) ready")

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
