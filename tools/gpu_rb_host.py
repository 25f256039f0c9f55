#!/usr/bin/env python
"""Is the launch loop host-bound?  Host time to enqueue vs device time to run n launches of the fused block."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from image_super_resolution_b200 import ops
from image_super_resolution_b200._lib import Context, lib, check
dev = torch.device("cuda:0"); ctx = Context.get(0)
g = torch.Generator().manual_seed(0)
rb = ops.ResBlock(ctx, torch.randn(32, 64, 3, 3, generator=g) / 24, torch.zeros(32), torch.randn(64, 32, 3, 3, generator=g) / 17,
                  torch.zeros(64), "fp16")
N, H, W = 256, 64, 64
x = torch.randn(N, 8, H, W, 8, device=dev).half(); y = torch.empty_like(x)
def dev_time(n=5):
    for _ in range(2): rb.forward(x, y, 0.2)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): rb.forward(x, y, 0.2)
    b.record(); torch.cuda.synchronize()
    return 1e3 * a.elapsed_time(b) / n
ctx.debug_set(3, 99)        # 100 launches per call, enqueued from C
for skip, name in ((1024, "full"), (0, "full, PDL"), (1024 + 7, "protocol only"), (7, "protocol only, PDL"), (1024 + 2048, "prologue only"), (2048, "prologue only, PDL"), (1024 + 4096, "immediate return"), (4096, "immediate return, PDL")):
    ctx.debug_set(2, skip)
    print(f"skip={skip:5d} ({name}): {dev_time() / 100:.2f} us per launch")
ctx.debug_set(2, 1024 + 4096)
for thr, sm in ((672, 0), (128, -1), (128, 0), (672, 100 * 1024)):
    ctx.debug_set(4, thr); ctx.debug_set(5, sm)
    print(f"immediate return, threads {thr}, smem {sm}: {dev_time() / 100:.2f} us per launch")
ctx.debug_set(2, 0); ctx.debug_set(3, 0); ctx.debug_set(4, 0); ctx.debug_set(5, -1)

# GPU-side timeline of 20 launches enqueued from C: %globaltimer of the earliest CTA start / latest CTA end per launch
ctx.debug_set(3, 19)
for skip, name in ((1024, "full"), (0, "full, PDL"), (1024 + 7, "protocol only"), (1024 + 2, "no MMA"), (1024 + 4, "no global I/O"), (1024 + 2048, "prologue only"), (1024 + 4096, "immediate return")):
    ctx.debug_set(2, skip)
    rb.forward(x, y, 0.2); torch.cuda.synchronize()
    tr = torch.zeros(4096, dtype=torch.int64, device=dev)
    rec = tr[2304:2304 + 1600].view(400, 4); rec[:, 0] = 1 << 62; rec[:, 2] = 1 << 62
    check(lib().srb_debug_trace(ctx.handle, tr.data_ptr()))
    rb.forward(x, y, 0.2); torch.cuda.synchronize()
    check(lib().srb_debug_trace(ctx.handle, None))
    r = tr.cpu()[2304:2304 + 80].view(20, 4)
    t0 = int(r[0, 0])
    rows = []
    for k in range(8, 14):
        rows.append(f"[start {int(r[k,0])-t0} (+{int(r[k,1]-r[k,0])}) first end +{int(r[k,2]-r[k,0])} last end +{int(r[k,3]-r[k,0])} | gap to next start {int(r[k+1,0]-r[k,3])}]")
    print(f"skip={skip} ({name}):\n  " + "\n  ".join(rows))
ctx.debug_set(2, 0); ctx.debug_set(3, 0)
