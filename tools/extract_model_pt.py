#!/usr/bin/env python
"""Extract the live fp32 weights of the reference's TorchScript archive into a plain .npz.

The reference ships `model.pt` as a `torch.jit.trace` archive (reference utils/models.py:801-802,
opened by rs.py:48).  /root/reference does not exist on the GPU box, so the weights (data, not
source) are extracted ONCE here into `tests/golden/model_pt_weights.npz`; the product loader
(`image_super_resolution_b200.load`) accepts either the original archive or this npz.

Dead `store_bn.*` BatchNorm tensors (never read by the traced forward) are dropped.

usage: python tools/extract_model_pt.py [/root/reference/model.pt] [out.npz]
"""
import sys
import numpy as np
import torch


def main():
    src = sys.argv[1] if len(sys.argv) > 1 else "/root/reference/model.pt"
    dst = sys.argv[2] if len(sys.argv) > 2 else "tests/golden/model_pt_weights.npz"
    m = torch.jit.load(src, "cpu").eval()
    out = {}
    for k, v in m.state_dict().items():
        if "store_bn" in k:
            continue
        out[k] = v.detach().cpu().numpy()
    np.savez(dst, **out)
    n = sum(int(v.size) for v in out.values())
    print(f"wrote {dst}: {len(out)} tensors, {n} values")


if __name__ == "__main__":
    main()
