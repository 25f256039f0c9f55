"""ORACLE (test infrastructure, NOT product code) — CPU fp32 restatement of the `model.pt` graph.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` legs may
import this module.  The product path (`image_super_resolution_b200`) never does.

What it restates (reference file:line; archive = /root/reference/model.pt, a torch.jit.trace):
  * Normalize.forward           utils/datasets.py:65-71  (traced with a uint8 example, so the
                                `/max_pixel_value` branch is baked in unconditionally)
  * Conv.fuseforward            utils/models.py:103-106  act(conv(x)+b), BN already folded (:366-406)
  * ResidualBlock1.forward      utils/models.py:208-209  x + m(x), m = Conv(64,32,3,LReLU.2), Conv(32,64,3)
  * Denoise.forward (old rev.)  archive __torch__.py: conv0 -> 8 blocks -> conv1 -> add(conv0_out) -> conv2
  * ConvWithoutBN + Tanh        utils/models.py:195-196
  * TanhToArrayImage.forward    utils/models.py:448-451  round((t+1)/2*255) half-to-even -> uint8

The arithmetic of the reference lives in PyTorch ATen (aten::_convolution -> oneDNN on CPU);
the oracle calls the same ATen ops in fp32 on CPU, in the same order, from the extracted weights
(`tests/golden/model_pt_weights.npz`, made by tools/extract_model_pt.py).

Parity pinning: the reference has NO tests or golden vectors.  This oracle is pinned
against outputs of the reference itself run in the build container: `tools/make_golden.py` executes
`torch.jit.load('/root/reference/model.pt')` on seeded inputs and stores input/output pairs under
`tests/golden/`; `tests/test_oracle.py` checks this restatement against them bit-for-bit (0 LSB).
"""
from __future__ import annotations

import os
import numpy as np
import torch
import torch.nn.functional as F

_HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_WEIGHTS = os.path.join(_HERE, "..", "tests", "golden", "model_pt_weights.npz")

N_BLOCKS = 8
LRELU_SLOPE = 0.2
RF_RADIUS = 25  # 4 (9x9) + 8*2*1 (3x3) + 1 (3x3) + 4 (9x9), SURVEY F6
FLOP_PER_PX = 725760


def load_weights(path: str | None = None) -> dict[str, torch.Tensor]:
    """state_dict of model.pt (minus dead store_bn tensors) as fp32 CPU tensors."""
    z = np.load(path or DEFAULT_WEIGHTS)
    return {k: torch.from_numpy(z[k].copy()) for k in z.files}


def normalize_u8(x: torch.Tensor, sd: dict) -> torch.Tensor:
    """utils/datasets.py:65-71 as traced: to(f32); /=255; -=mean; /=std (three separate roundings)."""
    x = x.to(torch.float32)
    x = x / sd["net.0.max_pixel_value"]
    x = x - sd["net.0.mean"]
    x = x / sd["net.0.std"]
    return x


def tanh_to_u8(t: torch.Tensor, sd: dict) -> torch.Tensor:
    """utils/models.py:448-451."""
    y = (t + 1.0) / 2.0
    y = y * sd["net.2.max_pixel_value"]
    return y.round().to(torch.uint8)


def denoise_forward(x: torch.Tensor, sd: dict, intermediates: dict | None = None) -> torch.Tensor:
    """Old-revision Denoise graph inside model.pt on a normalised float NCHW tensor -> tanh output."""
    p = "net.1."
    x0 = F.leaky_relu(F.conv2d(x, sd[p + "conv0.0.conv.weight"], sd[p + "conv0.0.conv.bias"], padding=4),
                      LRELU_SLOPE)
    if intermediates is not None:
        intermediates["conv0"] = x0
    r = x0
    for i in range(N_BLOCKS):
        q = f"{p}residual.{i}.m."
        h = F.leaky_relu(F.conv2d(r, sd[q + "0.conv.weight"], sd[q + "0.conv.bias"], padding=1), LRELU_SLOPE)
        r = r + F.conv2d(h, sd[q + "1.conv.weight"], sd[q + "1.conv.bias"], padding=1)
        if intermediates is not None:
            intermediates[f"block{i}.hidden"] = h
            intermediates[f"block{i}"] = r
    c1 = F.conv2d(r, sd[p + "conv1.conv.weight"], sd[p + "conv1.conv.bias"], padding=1)
    s = x0 + c1
    if intermediates is not None:
        intermediates["skip"] = s
    t = torch.tanh(F.conv2d(s, sd[p + "conv2.0.conv.weight"], sd[p + "conv2.0.conv.bias"], padding=4))
    if intermediates is not None:
        intermediates["tanh"] = t
    return t


@torch.no_grad()
def forward_u8(x_u8: torch.Tensor, sd: dict | None = None, intermediates: dict | None = None) -> torch.Tensor:
    """Whole model.pt graph: uint8 NCHW [N,3,H,W] -> uint8 NCHW (Model.forward, utils/models.py:738-739)."""
    sd = sd or load_weights()
    assert x_u8.dtype == torch.uint8 and x_u8.dim() == 4 and x_u8.shape[1] == 3
    return tanh_to_u8(denoise_forward(normalize_u8(x_u8, sd), sd, intermediates), sd)


@torch.no_grad()
def forward_tanh(x_u8: torch.Tensor, sd: dict | None = None) -> torch.Tensor:
    """Pre-rounding float output (tanh in [-1,1]) of the graph for tolerance studies."""
    sd = sd or load_weights()
    return denoise_forward(normalize_u8(x_u8, sd), sd)


# ---------------------------------------------------------------------------------------------
# rs.py tiling semantics (rs.py:16-27 sliding_window, :90-111 stitch) and exact halo tiling.
# ---------------------------------------------------------------------------------------------
def sliding_window(image: torch.Tensor, step, windowSize=None):
    """Restatement of rs.py:16-27: non-overlapping windows, ragged last row/col; yields (step,x,y,view)."""
    if windowSize is None:
        windowSize = step
    if isinstance(step, int):
        step = [step] * 2
    step = list(step)
    step[0] = min(image.shape[-2], step[0])
    step[1] = min(image.shape[-1], step[1])
    if isinstance(windowSize, int):
        wh, ww = windowSize, windowSize
    else:
        wh, ww = windowSize
    for y in range(0, image.shape[-2], step[0]):
        for x in range(0, image.shape[-1], step[1]):
            yield step, x, y, image[..., y:y + wh, x:x + ww]


@torch.no_grad()
def rs_image_reference(image_u8: torch.Tensor, window: int = 96, sd: dict | None = None) -> torch.Tensor:
    """rs.py:82-111 still-image path: each non-overlapping window forwarded on its own (batch 1,
    independently zero padded by the convs), results pasted in raster order.  [3,H,W] u8 -> [3,H,W] u8."""
    sd = sd or load_weights()
    out = torch.zeros_like(image_u8)
    for _, x, y, win in sliding_window(image_u8, window):
        o = forward_u8(win.unsqueeze(0), sd)[0]
        out[..., y:y + o.shape[-2], x:x + o.shape[-1]] = o
    return out


@torch.no_grad()
def halo_tiled(image_u8: torch.Tensor, tile_h: int, tile_w: int, halo: int = RF_RADIUS,
               sd: dict | None = None) -> torch.Tensor:
    """Exact tiling: each tile extended by the receptive-field radius, clipped at the image border
    (so border zero padding is identical to the whole-image forward); interior pasted back."""
    sd = sd or load_weights()
    _, H, W = image_u8.shape
    out = torch.zeros_like(image_u8)
    for y in range(0, H, tile_h):
        for x in range(0, W, tile_w):
            y0, x0 = max(0, y - halo), max(0, x - halo)
            y1, x1 = min(H, y + tile_h + halo), min(W, x + tile_w + halo)
            o = forward_u8(image_u8[:, y0:y1, x0:x1].unsqueeze(0), sd)[0]
            th, tw = min(tile_h, H - y), min(tile_w, W - x)
            out[:, y:y + th, x:x + tw] = o[:, y - y0:y - y0 + th, x - x0:x - x0 + tw]
    return out


def psnr(a: torch.Tensor, b: torch.Tensor, peak: float = 255.0) -> float:
    mse = torch.mean((a.to(torch.float64) - b.to(torch.float64)) ** 2).item()
    return float("inf") if mse == 0 else 10.0 * float(np.log10(peak * peak / mse))
