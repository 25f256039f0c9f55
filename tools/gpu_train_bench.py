#!/usr/bin/env python
"""Training-step timing of the drop-in EResNet (BASELINE configs[4] shape): samples/s, host vs device time."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from image_super_resolution_b200 import models
from image_super_resolution_b200.training import L1Loss

nb = int(sys.argv[1]) if len(sys.argv) > 1 else 16
B = int(sys.argv[2]) if len(sys.argv) > 2 else 16
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 10
dev = torch.device("cuda:0")
torch.manual_seed(100)
net = models.EResNet(nb, 0.2, 2).to(dev).train()
g = torch.Generator().manual_seed(100)
hr = (torch.rand(B, 3, 128, 128, generator=g) * 2 - 1).to(dev)
lr = torch.randn(B, 3, 64, 64, generator=g).to(dev)
loss_fn = L1Loss().to(dev)
opt = torch.optim.Adam(net.parameters(), lr=1e-4)
scaler = torch.amp.GradScaler("cuda")

def step():
    opt.zero_grad()
    with torch.autocast(device_type="cuda"):
        loss = loss_fn(net(lr), hr)
    scaler.scale(loss).backward()
    scaler.unscale_(opt)
    torch.nn.utils.clip_grad_norm_(net.parameters(), 10)
    scaler.step(opt)
    scaler.update()
    return loss

for _ in range(3):
    step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter(); e0.record()
for _ in range(steps):
    l = step()
e1.record(); th = time.perf_counter() - t0
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / steps
print(f"EResNet({nb}) B={B}: {ms:.2f} ms/step  {B / ms * 1e3:.1f} samples/s  host-issue {th / steps * 1e3:.2f} ms/step  loss {l.item():.4f}")
# phase split
def timed(fn, n=5):
    torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t = time.perf_counter(); a.record()
    for _ in range(n): fn()
    b.record(); h = time.perf_counter() - t; torch.cuda.synchronize()
    return a.elapsed_time(b) / n, h / n * 1e3
plan = net._train_plans[dev]; plan.use_graphs = False
print("repack   dev/host ms:", timed(plan.repack))
with torch.no_grad():
    print("forward  dev/host ms:", timed(lambda: plan.forward(lr)))
gout = torch.randn(B, 3, 128, 128, device=dev) * 1e-3
def bwd():
    plan.forward(lr); plan.backward(gout)
print("fwd+bwd  dev/host ms:", timed(bwd))
flop = 289.0e9 * B * nb / 16
print(f"fwd+bwd algorithmic {flop / 1e12:.2f} TFLOP -> {flop / (timed(bwd)[0] * 1e-3) / 1e12:.1f} TFLOP/s")
