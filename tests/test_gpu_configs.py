"""GPU parity at the shapes BASELINE.json's configs are quoted on (only config 2 was covered).

  config 1   repo model.pt via rs.py on one 256 x 256 RGB image (window 96: 9 windows of 96 / 96 / 64, batch 1) — against
             a golden produced by the reference's own rs.sliding_window loop on the reference's model.pt (CPU fp32);
  config 3   one 1080 x 1920 frame (17 overlapping strips per row in the row-fold kernels) against the fp32 oracle;
  config 4   a 4096 x 4096 image: a 1024 x 1024 crop against the fp32 oracle run on the crop + its 25-px halo, and the
             2 x 4 grid of 2048 x 1024 halo tiles bit-equal to the whole-image forward;
  config 5   EResNet(16, 0.2, 2) — forward against the reference class's golden output, and one full training step
             (B = 2, LR 64 x 64) against fp32 autograd of the oracle — in tests/test_nets.py / tests/test_gpu_train.py.

Tolerance (SURVEY §8(c), fp16 operands / fp32 accumulate): max-abs <= 2 LSB on uint8 pixels, >= 97 % of pixels exact.
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WEIGHTS = os.path.join(ROOT, "tests", "golden", "model_pt_weights.npz")


@pytest.fixture(scope="module")
def model():
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    import image_super_resolution_b200 as srb
    return srb.load(WEIGHTS).to("cuda:0").eval()


def _close(y, ref):
    d = (y.int() - ref.int()).abs()
    return int(d.max()), float((d == 0).float().mean())


def test_config1_rs_py_256x256_window96(model, golden_dir):
    """BASELINE configs[0] through the rs.py drop-in: the reference's seams (tile_mode="reference") and the whole image."""
    from image_super_resolution_b200 import rs
    g = np.load(os.path.join(golden_dir, "rs_config1_golden.npz"))
    img = torch.from_numpy(g["image"])
    assert tuple(img.shape) == (3, 256, 256)
    out = rs.upscale_image(model, img, 96, "reference")
    mx, exact = _close(out, torch.from_numpy(g["rs_window96"]))
    assert mx <= 2 and exact >= 0.97, (mx, exact)
    one_by_one = torch.zeros_like(img)                       # and as rs.py:90-111 issues it: one forward per window, batch 1
    for _, x, y, win in rs.sliding_window(img, 96):
        o = model(win.unsqueeze(0).contiguous().cuda())[0].cpu()
        one_by_one[:, y:y + o.shape[1], x:x + o.shape[2]] = o
    assert torch.equal(one_by_one, out)                      # batching equal-shaped windows changes nothing
    whole = rs.upscale_image(model, img, 96, "whole")
    mx, exact = _close(whole, torch.from_numpy(g["whole"]))
    assert mx <= 2 and exact >= 0.97, (mx, exact)
    halo = rs.upscale_image(model, img, 96, "halo", tile=(96, 96))
    assert torch.equal(halo, whole)


def test_config3_one_1080p_frame_vs_oracle(model):
    from oracle import modelpt_oracle as mo
    g = torch.Generator().manual_seed(100)
    lo = torch.rand(1, 3, 135, 240, generator=g)             # SURVEY §8(d) "smooth + noise" frame
    img = torch.nn.functional.interpolate(lo, size=(1080, 1920), mode="bicubic", align_corners=False).clamp(0, 1)
    x = ((img + 0.03 * torch.randn(1, 3, 1080, 1920, generator=g)).clamp(0, 1) * 255).round().to(torch.uint8)
    y = model(x.cuda()).cpu()
    torch.set_num_threads(os.cpu_count() or 1)
    ref = mo.forward_u8(x)
    mx, exact = _close(y, ref)
    print(f"1080p frame: max |diff| {mx} LSB, exact {exact:.4f}")
    assert mx <= 2 and exact >= 0.97, (mx, exact)
    # the video path's store (BGR, HWC) holds the same bytes
    f = model(x.cuda(), layout="hwc_bgr").cpu()
    assert torch.equal(f[0], y[0].flip(0).permute(1, 2, 0))


def test_config4_4096_image_crop_vs_oracle_and_halo_grid(model):
    from oracle import modelpt_oracle as mo
    from image_super_resolution_b200 import rs
    g = torch.Generator().manual_seed(100)
    img = torch.randint(0, 256, (3, 4096, 4096), dtype=torch.uint8, generator=g)
    whole = model(img.unsqueeze(0).cuda())[0].cpu()
    R = mo.RF_RADIUS
    y0, x0, S = 1500, 2100, 1024                              # interior crop: its halo lies inside the image
    torch.set_num_threads(os.cpu_count() or 1)
    ref = mo.forward_u8(img[:, y0 - R:y0 + S + R, x0 - R:x0 + S + R].unsqueeze(0))[0][:, R:R + S, R:R + S]
    mx, exact = _close(whole[:, y0:y0 + S, x0:x0 + S], ref)
    print(f"4096^2 image, 1024^2 crop: max |diff| {mx} LSB, exact {exact:.4f}")
    assert mx <= 2 and exact >= 0.97, (mx, exact)
    # a corner crop (zero padding on two sides) as well
    ref = mo.forward_u8(img[:, :256 + R, 4096 - 256 - R:].unsqueeze(0))[0][:, :256, R:]
    mx, exact = _close(whole[:, :256, 4096 - 256:], ref)
    assert mx <= 2 and exact >= 0.97, (mx, exact)
    # a 2 x 4 grid of 2048 x 1024 tiles, each with its 25-px halo, equals the whole image bit for bit
    tiles = rs.upscale_image(model, img, 96, "halo", tile=(2048, 1024))
    assert torch.equal(tiles, whole)
    del whole, tiles
    torch.cuda.empty_cache()
