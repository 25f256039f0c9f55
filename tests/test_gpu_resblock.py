"""GPU parity of the fused residual-block kernel (srb_resblock.cuh) against fp32 torch on the same fp16-rounded operands
(hidden activation rounded to fp16 like the unfused layer chain).  Tolerance: |err| <= 4e-3 * max|ref| + 2e-3."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    return torch.device("cuda:0")


def _case(dev, N, H, W, seed=0, grid=0):
    from image_super_resolution_b200 import ops
    from image_super_resolution_b200._lib import Context
    ctx = Context.get(0)
    g = torch.Generator().manual_seed(seed)
    wA = torch.randn(32, 64, 3, 3, generator=g) / 24
    bA = torch.randn(32, generator=g) * 0.1
    wB = torch.randn(64, 32, 3, 3, generator=g) / 17
    bB = torch.randn(64, generator=g) * 0.1
    x = torch.randn(N, 64, H, W, generator=g)
    q = lambda t: t.half().float()  # noqa: E731
    hid = q(F.leaky_relu(F.conv2d(q(x), q(wA), bA, padding=1), 0.2))
    ref = q(x) + F.conv2d(hid, q(wB), bB, padding=1)
    rb = ops.ResBlock(ctx, wA, bA, wB, bB, "fp16")
    xc = ops.nchw_to_c8(x.to(dev), dtype="fp16")
    yc = torch.full_like(xc, float("nan"))
    if grid:
        ctx.set_option(1, grid)
    try:
        rb.forward(xc, yc, 0.2)
        torch.cuda.synchronize()
        ctx.check_device()
    finally:
        ctx.set_option(1, 0)
    out = ops.c8_to_nchw(yc, 64).cpu()
    err = (out - ref).abs().max().item()
    assert np.isfinite(err) and err <= 4e-3 * ref.abs().max().item() + 2e-3, err


@pytest.mark.parametrize("N,H,W", [(2, 12, 64), (3, 20, 64), (1, 5, 8), (5, 9, 16), (2, 33, 128), (3, 17, 60), (2, 1, 64), (7, 2, 28)])
def test_resblock_matches_fp32_reference(dev, N, H, W):
    _case(dev, N, H, W, seed=N * 100 + H)


def test_resblock_many_rows_per_cta_and_segment_splits(dev):
    """few CTAs (long marches, ring wrap-arounds) and many CTAs (segments that split images mid-way)"""
    _case(dev, 6, 64, 64, seed=1, grid=2)
    _case(dev, 6, 64, 64, seed=2, grid=0)
    _case(dev, 40, 64, 64, seed=3, grid=0)


# widths that do not divide a 128-pixel tile (idle slots), and rows wider than a tile (overlapping strips: the result of a
# strip's edge groups comes from the neighbouring strip), ragged ends included
ANY_WIDTH = [(2, 9, 56), (3, 7, 24), (1, 7, 96), (2, 6, 100), (1, 6, 136), (2, 5, 250), (1, 4, 360), (1, 3, 129), (2, 4, 240)]


@pytest.mark.parametrize("N,H,W", ANY_WIDTH)
def test_resblock_any_width(dev, N, H, W):
    _case(dev, N, H, W, seed=N * 100 + H + W)


def test_resblock_wide_rows_long_marches(dev):
    _case(dev, 1, 70, 300, seed=11, grid=3)
    _case(dev, 2, 40, 200, seed=12, grid=0)


def test_resblock_supported_every_width(dev):
    from image_super_resolution_b200 import ops
    assert ops.ResBlock.supported(64, 64) and ops.ResBlock.supported(7, 128) and ops.ResBlock.supported(64, 96)
    assert ops.ResBlock.supported(1080, 1920) and not ops.ResBlock.supported(0, 64)


# ---- row-fold single conv (srb_rowconv.cuh): conv1 + long skip of the model.pt graph ----
def _rowconv_case(dev, N, H, W, seed=0, grid=0, res=True, lrelu=False):
    from image_super_resolution_b200 import ops
    from image_super_resolution_b200._lib import Context
    ctx = Context.get(0)
    g = torch.Generator().manual_seed(seed)
    w = torch.randn(64, 64, 3, 3, generator=g) / 24
    b = torch.randn(64, generator=g) * 0.1
    x = torch.randn(N, 64, H, W, generator=g)
    r = torch.randn(N, 64, H, W, generator=g)
    q = lambda t: t.half().float()  # noqa: E731
    ref = F.conv2d(q(x), q(w), b, padding=1)
    if lrelu:
        ref = F.leaky_relu(ref, 0.2)
    if res:
        ref = ref + q(r)
    rc = ops.RowConv(ctx, w, b, "fp16")
    xc = ops.nchw_to_c8(x.to(dev), dtype="fp16")
    rcc = ops.nchw_to_c8(r.to(dev), dtype="fp16") if res else None
    yc = torch.full_like(xc, float("nan"))
    if grid:
        ctx.set_option(1, grid)
    try:
        rc.forward(xc, yc, rcc, slope=0.2, lrelu=lrelu)
        torch.cuda.synchronize()
        ctx.check_device()
    finally:
        ctx.set_option(1, 0)
    out = ops.c8_to_nchw(yc, 64).cpu()
    err = (out - ref).abs().max().item()
    assert np.isfinite(err) and err <= 4e-3 * ref.abs().max().item() + 2e-3, err


@pytest.mark.parametrize("N,H,W", [(2, 12, 64), (3, 20, 64), (1, 5, 8), (5, 9, 16), (2, 33, 128), (3, 17, 60), (2, 1, 64), (7, 2, 28)])
def test_rowconv_matches_fp32_reference(dev, N, H, W):
    _rowconv_case(dev, N, H, W, seed=N * 100 + H)


@pytest.mark.parametrize("N,H,W", ANY_WIDTH)
def test_rowconv_any_width(dev, N, H, W):
    _rowconv_case(dev, N, H, W, seed=N * 100 + H + W)


def test_rowconv_variants_and_segment_splits(dev):
    _rowconv_case(dev, 6, 64, 64, seed=1, grid=2)
    _rowconv_case(dev, 6, 64, 64, seed=2, grid=0)
    _rowconv_case(dev, 3, 19, 32, seed=3, res=False)
    _rowconv_case(dev, 3, 19, 32, seed=4, res=True, lrelu=True)
    _rowconv_case(dev, 9, 40, 64, seed=5, grid=37)


# ---- row-fold 9x9 tail conv (srb_tail.cuh): conv2 + tanh -> uint8 of the model.pt graph ----
def _tail_case(dev, N, H, W, seed=0, grid=0):
    from image_super_resolution_b200 import ops
    from image_super_resolution_b200._lib import Context
    ctx = Context.get(0)
    g = torch.Generator().manual_seed(seed)
    w = torch.randn(3, 64, 9, 9, generator=g) / 72
    b = torch.randn(3, generator=g) * 0.1
    x = torch.randn(N, 64, H, W, generator=g)
    q = lambda t: t.half().float()  # noqa: E731
    pre = F.conv2d(q(x), q(w), b, padding=4)
    t = ops.Tail9(ctx, w, b, "fp16")
    xc = ops.nchw_to_c8(x.to(dev), dtype="fp16")
    if grid:
        ctx.set_option(1, grid)
    try:
        yf = t.forward(xc, tanh=False, u8=False, out=torch.full((N, 3, H, W), float("nan"), device=dev))
        yu = t.forward(xc, tanh=True, u8=True, out=torch.full((N, 3, H, W), 77, dtype=torch.uint8, device=dev))
        torch.cuda.synchronize()
        ctx.check_device()
    finally:
        ctx.set_option(1, 0)
    err = (yf.cpu() - pre).abs().max().item()
    assert np.isfinite(err) and err <= 2e-3 * pre.abs().max().item() + 1e-3, err
    want = torch.round((torch.tanh(pre) + 1) / 2 * 255).to(torch.int16)
    d = (yu.cpu().to(torch.int16) - want).abs()
    assert d.max().item() <= 1 and (d == 0).float().mean().item() > 0.98, (d.max().item(), (d == 0).float().mean().item())


@pytest.mark.parametrize("N,H,W", [(2, 12, 64), (3, 20, 64), (1, 5, 8), (5, 9, 16), (2, 33, 128), (3, 17, 60), (2, 1, 64), (7, 2, 28), (1, 3, 32)])
def test_tail9_matches_fp32_reference(dev, N, H, W):
    _tail_case(dev, N, H, W, seed=N * 100 + H)


@pytest.mark.parametrize("N,H,W", ANY_WIDTH)
def test_tail9_any_width(dev, N, H, W):
    _tail_case(dev, N, H, W, seed=N * 100 + H + W)


def test_tail9_wide_rows_long_marches(dev):
    _tail_case(dev, 1, 70, 300, seed=11, grid=3)
    _tail_case(dev, 2, 40, 200, seed=12, grid=0)


def test_tail9_long_marches_and_segment_splits(dev):
    _tail_case(dev, 6, 64, 64, seed=1, grid=2)
    _tail_case(dev, 6, 64, 64, seed=2, grid=0)
    _tail_case(dev, 9, 40, 64, seed=5, grid=37)
