#!/usr/bin/env python
"""Data-parallel check on real GPUs (torchrun, NCCL): after one backward on per-rank shards, every rank's gradient arena
equals the single-process gradient of the global batch.  Usage: torchrun --nproc-per-node 2 tools/ddp_check.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
from image_super_resolution_b200 import models, train as T
from image_super_resolution_b200.training import L1Loss

rank, world, dev = T.init_distributed()
torch.manual_seed(100 + rank)
net = models.EResNet(2, 0.2, 2).to(dev).train()
T.broadcast_parameters(net)
g = torch.Generator().manual_seed(5)
hr = torch.rand(4 * world, 3, 64, 64, generator=g) * 2 - 1
lr = torch.randn(4 * world, 3, 32, 32, generator=g)
loss_fn = L1Loss().to(dev)
loss_fn(net(T.shard_of(lr, rank, world).to(dev)), T.shard_of(hr, rank, world).to(dev)).backward()
plan = net._train_plans[dev]
got = plan.flat_grad.clone()
# single-process reference on the global batch (all-reduce switched off)
for p in net.parameters():
    p.grad = None
plan.allreduce = False
loss_fn(net(lr.to(dev)), hr.to(dev)).backward()
want = plan.flat_grad.clone()
cos = float(torch.dot(got.double(), want.double()) / (got.double().norm() * want.double().norm()))
rel = float((got - want).norm() / want.norm())
all_ok = torch.tensor([1 if (cos > 0.9999 and rel < 2e-2) else 0], device=dev)
if world > 1:
    dist.all_reduce(all_ok, op=dist.ReduceOp.MIN)
print(f"rank {rank}/{world}: cosine {cos:.6f} rel-err {rel:.2e} ok={int(all_ok.item())}", flush=True)
if world > 1:
    dist.destroy_process_group()
sys.exit(0 if int(all_ok.item()) else 1)
