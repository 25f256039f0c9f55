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
tr = torch.zeros(1024, dtype=torch.int64, device=dev)
for i in range(3):
    if i == 2: check(lib().srb_debug_trace(ctx.handle, tr.data_ptr()))
    rb.forward(x, y, 0.2)
torch.cuda.synchronize()
check(lib().srb_debug_trace(ctx.handle, None)); ctx.debug_set(2, 0)
t = tr.cpu().view(64, 16)
t0 = int(t[0, 0])
print(f"skip={skip}: cycles relative to the first MMA iteration")
print("idx | mma: ready issued (len, period) | hid-epi: wait  full  done (len) | out-epi: wait  full  done (len)")
pm = None
for i in range(40):
    r = [int(v) - t0 if int(v) else 0 for v in t[i]]
    per = r[0] - pm if pm is not None else 0
    pm = r[0]
    print(f"      hid: ld {r[11]-r[3]:5d} hsempty {r[12]-r[11]:5d} sts {r[13]-r[12]:5d} stwait {r[14]-r[13]:5d} fence+arrive {r[4]-r[14]:5d} | out: ld {r[8]-r[6]:5d} math+stg {r[9]-r[8]:5d} stwait {r[10]-r[9]:5d} arrive {r[7]-r[10]:5d} | half0: math {r[15]-r[8]:5d}")
    print(f"{i:3d} | {r[0]:7d} {r[1]:7d} ({r[1]-r[0]:5d}, {per:5d}) | {r[2]:7d} {r[3]:7d} {r[4]:7d} ({r[4]-r[3]:5d}) | {r[5]:7d} {r[6]:7d} {r[7]:7d} ({r[7]-r[6]:5d})")
