#!/usr/bin/env python
"""Per-phase time of the drop-in train() step (host + device, synchronised after every phase) and the un-synchronised step."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from torch.nn.utils import clip_grad_norm_
from image_super_resolution_b200 import models
from image_super_resolution_b200 import train as T
from image_super_resolution_b200.training import L1Loss
dev = torch.device("cuda:0")
B = 16
torch.manual_seed(100)
opts = dict(kv.split("=") for kv in sys.argv[2:])
net = getattr(models, opts.pop("cls", "EResNet"))(16, 0.2, 2).to(dev)        # cls=ResNet: train-mode BatchNorm (BNTrainPlan)
loss_name = opts.pop("loss", "l1")
for k, v in opts.items():                    # e.g. train_conv_fold=1 train_chain_nt4=1
    setattr(net, k, int(v))
g = torch.Generator().manual_seed(100)
hr = (torch.rand(B, 3, 128, 128, generator=g) * 2 - 1).to(dev)
lr = torch.randn(B, 3, 64, 64, generator=g).to(dev)
from image_super_resolution_b200.training import MSELoss
loss_fn = (MSELoss() if loss_name == "mse" else L1Loss()).to(dev)
mode = sys.argv[1] if len(sys.argv) > 1 else "foreach"      # foreach | fused (torch Adam fused=True) | flat (optim.FlatAdam tail)
fused = mode
if mode == "flat":
    opt = T.FlatAdam(net.parameters(), lr=1e-4, betas=(0.9, 0.999))
    scaler = T.GradScaler("cuda")
    ema = T.ModelEMA(net, tau=1000).attach(opt)
    clip_grad_norm_ = T.clip_grad_norm_
else:
    opt = torch.optim.Adam(net.parameters(), lr=1e-4, betas=(0.9, 0.999), fused=mode == "fused")
    scaler = torch.amp.GradScaler("cuda")
    ema = T.ModelEMA(net, tau=1000)
sched = torch.optim.lr_scheduler.LinearLR(opt, start_factor=1, end_factor=0.1, total_iters=1000)
T.train(net, ema, [(hr, lr)] * 3, loss_fn, opt, scaler, sched, 0, None)
acc = {}
def phase(name, fn):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    r = fn()
    torch.cuda.synchronize(); acc[name] = acc.get(name, 0.0) + (time.perf_counter() - t0) * 1e3
    return r
n = 10
for _ in range(n):
    phase("zero_grad", lambda: opt.zero_grad())
    def fwd():
        with torch.autocast(enabled=True, device_type="cuda"):
            preds = net(lr)
            return loss_fn(preds, hr)
    loss = phase("forward+loss", fwd)
    phase("backward", lambda: scaler.scale(loss).backward())
    phase("unscale_", lambda: scaler.unscale_(opt))
    phase("clip_grad_norm_", lambda: clip_grad_norm_(net.parameters(), 10))
    phase("scaler.step(Adam)", lambda: scaler.step(opt))
    phase("scaler.update", lambda: scaler.update())
    phase("schedule.step", lambda: sched.step())
    phase("ema.update", lambda: ema.update(net))
    phase("loss.item", lambda: loss.item())
print(f"Adam fused={fused}; per-phase ms (synchronised):")
for k, v in acc.items(): print(f"  {k:20s} {v / n:7.2f}")
print(f"  {'sum':20s} {sum(acc.values()) / n:7.2f}")
torch.cuda.synchronize(); t0 = time.perf_counter()
T.train(net, ema, [(hr, lr)] * n, loss_fn, opt, scaler, sched, 1, None)
torch.cuda.synchronize(); print(f"un-synchronised train(): {(time.perf_counter() - t0) * 1e3 / n:.2f} ms/step")
