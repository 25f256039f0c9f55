#!/usr/bin/env python
"""A few model.pt forwards at the bench shape (256 x 3 x 64 x 64 uint8), for ncu captures of every kernel of the step."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import image_super_resolution_b200 as srb

dev = torch.device("cuda:0")
model = srb.load(os.path.join(ROOT, "tests", "golden", "model_pt_weights.npz")).to(dev).eval()
g = torch.Generator().manual_seed(0)
x = torch.randint(0, 256, (256, 3, 64, 64), dtype=torch.uint8, generator=g).to(dev)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    y = model(x)
torch.cuda.synchronize()
print("ok", tuple(y.shape), int(y.sum()))
