import os, sys, time
sys.path.insert(0, "/root/repo")
import torch
from image_super_resolution_b200 import models, train as T
from image_super_resolution_b200.training import L1Loss
dev = torch.device("cuda:0"); B = 16
torch.manual_seed(100)
net = models.EResNet(16, 0.2, 2).to(dev)
g = torch.Generator().manual_seed(100)
hr = (torch.rand(B, 3, 128, 128, generator=g) * 2 - 1).to(dev); lr = torch.randn(B, 3, 64, 64, generator=g).to(dev)
loss_fn = L1Loss().to(dev)
opt = T.FlatAdam(net.parameters(), lr=1e-4); sc = T.GradScaler("cuda"); ema = T.ModelEMA(net, tau=1000).attach(opt)
sched = torch.optim.lr_scheduler.LinearLR(opt, start_factor=1, end_factor=0.1, total_iters=1000)
T.train(net, ema, [(hr, lr)] * 3, loss_fn, opt, sc, sched, 0, None)
torch.cuda.synchronize()
import cProfile, pstats
n = 20
t0 = time.perf_counter()
pr = cProfile.Profile(); pr.enable()
# host-only view: the loop body without the final read-back
model = net; model.train()
for i in range(n):
    opt.zero_grad()
    with torch.autocast(enabled=True, device_type="cuda"):
        loss = loss_fn(model(lr), hr)
    sc.scale(loss).backward()
    sc.unscale_(opt); T.clip_grad_norm_(model.parameters(), 10); sc.step(opt); sc.update(); sched.step(); ema.update(model)
pr.disable()
t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
print(f"host enqueue {1e3 * (t1 - t0) / n:.2f} ms/step, with device completion {1e3 * (t2 - t0) / n:.2f} ms/step")
pstats.Stats(pr).sort_stats("cumulative").print_stats(22)
