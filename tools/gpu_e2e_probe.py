#!/usr/bin/env python
"""Where the end-to-end (host batch -> host result) time of ModelPT.stream goes: fixed cost per call vs marginal cost per batch."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import image_super_resolution_b200 as srb

dev = torch.device("cuda:0")
model = srb.load(os.path.join(ROOT, "tests", "golden", "model_pt_weights.npz")).to(dev).eval()
x_host = torch.randint(0, 256, (256, 3, 64, 64), dtype=torch.uint8).pin_memory()
x = x_host.to(dev); y = torch.empty_like(x)


def batches(n):
    for _ in range(n):
        yield x_host


def run(K, depth=2):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n = 0
    for yh in model.stream(batches(K), depth=depth):
        n += 1
    t1 = time.perf_counter()
    assert n == K
    return (t1 - t0) * 1e3


for _ in range(2):
    run(5)
for depth in (1, 2, 3, 4):
    r = {K: min(run(K, depth) for _ in range(3)) for K in (1, 5, 20, 100)}
    marg = (r[100] - r[20]) / 80
    print(f"depth {depth}: " + "  ".join(f"K={K}: {v:7.3f} ms" for K, v in r.items()) + f"   marginal {marg:.4f} ms/batch, fixed {r[20] - 20 * marg:.3f} ms")
# device-resident reference and host enqueue cost
for _ in range(5):
    model.forward(x, out=y)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(100):
    model.forward(x, out=y)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"device-resident, 100 forwards back to back: host enqueue {(t1 - t0) * 1e4:.1f} us per forward, {(t2 - t0) * 1e4:.1f} us per forward to completion")
