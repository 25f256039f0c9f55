#!/usr/bin/env python
"""The fused optimizer tail on arenas of EResNet(16)'s size (11.7 M parameters): event-timed launches and achieved HBM
bandwidth (algorithmic bytes: grad_stats reads g = 4 B/param; adam_ema reads p, g, m, v, ema and writes p, m, v, ema =
36 B/param).  Also the command the ncu --set full capture of the two kernels is taken on."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from image_super_resolution_b200 import train as T

dev = torch.device("cuda:0")
n = 11_726_595
torch.manual_seed(0)
p = torch.nn.Parameter(torch.randn(n, device=dev) * 0.02)
opt = T.FlatAdam([p], lr=1e-4)
sc = T.GradScaler("cuda")
sc.scale(torch.zeros((), device=dev))
class _M(torch.nn.Module):
    def __init__(s):
        super().__init__(); s.p = p
ema = T.ModelEMA(_M(), tau=1000).attach(opt)
garena = torch.randn(opt.n, device=dev) * 65536.0
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
ts = {"grad_stats": [], "adam_ema": []}
for it in range(8):
    p.grad = garena[:n].view_as(p)
    flush.zero_()                                    # 256 MB > 126 MB L2: the arenas come from HBM
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    e[0].record(); sc.unscale_(opt); T.clip_grad_norm_([p], 10); e[1].record(); sc.step(opt); e[2].record()
    sc.update()
    ema.updates += 1
    torch.cuda.synchronize()
    if it >= 3:
        ts["grad_stats"].append(e[0].elapsed_time(e[1])); ts["adam_ema"].append(e[1].elapsed_time(e[2]))
for k, b in (("grad_stats", 4), ("adam_ema", 36)):
    ms = sorted(ts[k])[len(ts[k]) // 2]
    print(f"{k:11s} {ms * 1e3:7.1f} us  {b * n / ms / 1e6:7.1f} GB/s algorithmic ({b} B/param x {n} params)")
