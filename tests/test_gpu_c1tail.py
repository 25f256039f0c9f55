"""GPU tests of the fused conv1 + long skip + conv2 9x9 + tanh -> uint8 kernel (srb_c1tail.cuh): it must give, bit for bit,
what srb_rowconv_fwd followed by srb_tail9_fwd give (both are parity-tested against fp32 torch in tests/test_gpu_resblock.py) —
same K order per accumulator, same roundings; conv1's result only stops going through HBM.  Shapes cover two image rows per
tile (W = 64), ragged widths, a single image row per tile, wide rows cut into overlapping strips, runs of rows that start
and end inside an image (grids capped so that segment boundaries fall everywhere) and images shorter than the 9 x 9 halo."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WEIGHTS = os.path.join(ROOT, "tests", "golden", "model_pt_weights.npz")


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    return torch.device("cuda:0")


SHAPES = [(6, 64, 64, 0), (6, 64, 64, 2), (5, 40, 64, 3), (3, 33, 60, 0), (2, 17, 72, 0), (1, 50, 120, 5), (2, 24, 128, 0),
          (1, 70, 300, 3), (2, 40, 200, 0), (1, 30, 1000, 37), (3, 5, 64, 0), (1, 3, 136, 1), (40, 64, 64, 0), (7, 9, 88, 4)]


@pytest.mark.parametrize("N,H,W,grid", SHAPES)
@pytest.mark.parametrize("dtype", ["fp16", "bf16"])
def test_c1tail_equals_rowconv_then_tail(dev, N, H, W, grid, dtype):
    from image_super_resolution_b200 import ops
    from image_super_resolution_b200._lib import Context
    ctx = Context(0)
    g = torch.Generator().manual_seed(N * 100000 + H * 1000 + W)
    rc = ops.RowConv(ctx, torch.randn(64, 64, 3, 3, generator=g) / 24, torch.randn(64, generator=g) * 0.1, dtype)
    t9 = ops.Tail9(ctx, torch.randn(3, 64, 9, 9, generator=g) / 72, torch.randn(3, generator=g) * 0.1, dtype)
    tdt = torch.float16 if dtype == "fp16" else torch.bfloat16
    x = (torch.randn(N, 8, H, W, 8, generator=g) * 0.7).to(tdt).to(dev)
    skip = (torch.randn(N, 8, H, W, 8, generator=g) * 0.7).to(tdt).to(dev)
    assert ops.c1tail_supported(H, W)
    mid = torch.empty_like(x)
    rc.forward(x, mid, skip)
    want_u8 = t9.forward(mid)
    want_f32 = t9.forward(mid, u8=False)
    if grid:
        ctx.set_option(1, grid)
    got_u8 = ops.c1tail_forward(rc, t9, x, skip)
    got_f32 = ops.c1tail_forward(rc, t9, x, skip, u8=False)
    got_hwc = ops.c1tail_forward(rc, t9, x, skip, hwc_bgr=True)
    torch.cuda.synchronize()
    ctx.check_device()
    assert torch.equal(got_f32, want_f32), f"fp32: max diff {(got_f32 - want_f32).abs().max().item()}"
    assert torch.equal(got_u8, want_u8)
    assert torch.equal(got_hwc, want_u8.permute(0, 2, 3, 1).flip(-1))


def test_c1tail_refuses_narrow_rows(dev):
    from image_super_resolution_b200 import ops
    from image_super_resolution_b200._lib import Context
    ctx = Context(0)
    g = torch.Generator().manual_seed(1)
    rc = ops.RowConv(ctx, torch.randn(64, 64, 3, 3, generator=g) / 24, torch.zeros(64), "fp16")
    t9 = ops.Tail9(ctx, torch.randn(3, 64, 9, 9, generator=g) / 72, torch.zeros(3), "fp16")
    x = torch.zeros(1, 8, 16, 40, 8, dtype=torch.float16, device=dev)
    assert not ops.c1tail_supported(16, 40)
    with pytest.raises(RuntimeError):
        ops.c1tail_forward(rc, t9, x, x)


def test_model_pt_fused_and_split_tail_agree(dev):
    """The whole graph with the fused conv1 + tail launch against the two launches (SRB_PATH_SPLIT_C1TAIL): identical bytes."""
    import image_super_resolution_b200 as srb
    from image_super_resolution_b200 import _lib
    model = srb.load(WEIGHTS).to(dev).eval()
    g = torch.Generator().manual_seed(4)
    for shape in ((4, 3, 64, 64), (1, 3, 50, 136), (2, 3, 33, 60), (1, 3, 100, 300)):
        x = torch.randint(0, 256, shape, dtype=torch.uint8, generator=g).to(dev)
        assert model.launches_per_forward(shape[2], shape[3]) == 10
        y_fused = model(x).cpu()
        model._ctx.set_option(_lib.OPT_PATH, _lib.PATH_SPLIT_C1TAIL)
        try:
            y_split = model(x).cpu()
        finally:
            model._ctx.set_option(_lib.OPT_PATH, 0)
        assert torch.equal(y_fused, y_split), shape
    assert model.launches_per_forward(20, 40) == 11
    model._ctx.check_device()
