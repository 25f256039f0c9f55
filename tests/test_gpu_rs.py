"""GPU tests of the rs.py drop-in (tile modes, frame path)."""
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


def test_reference_tile_mode_reproduces_rs_py_seams(model, golden_dir):
    from image_super_resolution_b200 import rs
    g = np.load(os.path.join(golden_dir, "rs_tiling_golden.npz"))
    img = torch.from_numpy(g["image"])
    out = rs.upscale_image(model, img, 96, "reference")
    d = (out.int() - torch.from_numpy(g["rs_window96"]).int()).abs()
    assert d.max().item() <= 2 and (d == 0).float().mean().item() >= 0.97
    # and it is NOT the whole-image result: the seams of the reference are reproduced, not repaired
    dw = (out.int() - torch.from_numpy(g["whole"]).int()).abs()
    assert dw.max().item() > 10


def test_halo_tile_mode_equals_whole_image(model, golden_dir):
    from image_super_resolution_b200 import rs
    g = np.load(os.path.join(golden_dir, "rs_tiling_golden.npz"))
    img = torch.from_numpy(g["image"])
    whole = rs.upscale_image(model, img, 96, "whole")
    halo = rs.upscale_image(model, img, 96, "halo", tile=(64, 80))
    assert torch.equal(whole, halo)                       # exact, SURVEY F6
    d = (whole.int() - torch.from_numpy(g["whole"]).int()).abs()
    assert d.max().item() <= 2


@pytest.mark.parametrize("ndev,nb", [(1, 3), (2, 5), (2, 4)])
def test_sharded_model_and_frame_path(model, ndev, nb):
    """rs.upscale_frames: batch i on device i mod G through that replica's pipelined stream, frames leave in order as BGR HWC."""
    from image_super_resolution_b200 import rs
    if torch.cuda.device_count() < ndev:
        pytest.skip(f"needs {ndev} GPUs")
    sm = rs.ShardedModel(model, devices=list(range(ndev)))
    g = torch.Generator().manual_seed(3)
    batches = [torch.randint(0, 256, (2, 3, 40, 48), dtype=torch.uint8, generator=g) for _ in range(nb)]
    sink = rs.ByteSink(keep=True)
    n = rs.upscale_frames(sm, batches, sink)
    assert n == 2 * nb and sink.bytes == 2 * nb * 40 * 48 * 3
    for i in range(nb):
        want = model(batches[i].cuda())
        for j in range(2):
            assert np.array_equal(sink.kept[2 * i + j], want[j].flip(0).permute(1, 2, 0).cpu().numpy()), (i, j)
    sink_rgb = rs.ByteSink(keep=True)
    rs.upscale_frames(sm, batches[:2], sink_rgb, bgr=False)
    assert np.array_equal(sink_rgb.kept[3], model(batches[1].cuda())[1].permute(1, 2, 0).cpu().numpy())
    # HWC BGR frames in (a decoder's layout), HWC BGR frames out: the same bytes
    sink_hwc = rs.ByteSink(keep=True)
    rs.upscale_frames(sm, [b.permute(0, 2, 3, 1).flip(-1).contiguous() for b in batches], sink_hwc, in_layout="hwc_bgr")
    assert len(sink_hwc.kept) == 2 * nb and all(np.array_equal(a, b) for a, b in zip(sink_hwc.kept, sink.kept))


def test_stream_pipelined_host_batches_match_forward(model):
    """ModelPT.stream (copy / compute overlap, ring of device buffers) returns, in order, what forward() returns."""
    g = torch.Generator().manual_seed(7)
    batches = [torch.randint(0, 256, (3, 3, 40, 64), dtype=torch.uint8, generator=g) for _ in range(7)]
    batches.insert(3, torch.randint(0, 256, (2, 3, 24, 32), dtype=torch.uint8, generator=g))   # a shape change mid-stream
    want = [model(b.cuda()).cpu() for b in batches]
    got = [y.clone() for y in model.stream(iter(batches), depth=2)]
    assert len(got) == len(want)
    for a, b in zip(got, want):
        assert torch.equal(a, b)
    with pytest.raises(RuntimeError):
        list(model.stream([torch.zeros(1, 3, 8, 8)]))


@pytest.mark.parametrize("shape", [(2, 3, 40, 48), (1, 3, 24, 200), (3, 3, 64, 64), (1, 3, 37, 61)])
def test_hwc_bgr_store_equals_flip_permute_of_the_nchw_result(model, shape):
    """RGB2BGR + NCHW->HWC (rs.py:65-74, utils/datasets.py:113-124) folded into the tail conv's store is
    bit-identical to the glue applied to the NCHW result — direct call, pipelined host stream, and the tiled tail kernel."""
    from image_super_resolution_b200._lib import Context
    g = torch.Generator().manual_seed(13)
    x = torch.randint(0, 256, shape, dtype=torch.uint8, generator=g)
    want = model(x.cuda()).flip(1).permute(0, 2, 3, 1).contiguous().cpu()
    got = model.forward(x.cuda(), layout="hwc_bgr").cpu()
    assert got.shape == (shape[0], shape[2], shape[3], 3) and torch.equal(got, want)
    got_s = [y.clone() for y in model.stream(iter([x, x]), depth=2, layout="hwc_bgr")]
    assert all(torch.equal(y, want) for y in got_s)
    ctx = Context.get(0)
    ctx.set_option(2, 16384)                   # the tiled 9x9 tail kernel (fallback path) stores the same bytes
    try:
        got_t = model.forward(x.cuda(), layout="hwc_bgr").cpu()
    finally:
        ctx.set_option(2, 0)
    assert torch.equal(got_t, want)
    with pytest.raises(RuntimeError):
        model.forward(x, layout="hwc_bgr")    # host tensors go through stream()


@pytest.mark.parametrize("ndev", [1, 2])
def test_upscale_image_modes_agree_with_per_window_forwards(model, ndev):
    """rs.upscale_image with the device-resident canvas: "reference" (windows of equal shape batched) == one forward per
    rs.sliding_window window, pasted like rs.py:105-111; "halo" == "whole"; a ragged image, on one and on two devices."""
    from image_super_resolution_b200 import rs
    if torch.cuda.device_count() < ndev:
        pytest.skip(f"needs {ndev} GPUs")
    m = model if ndev == 1 else rs.ShardedModel(model, devices=list(range(ndev)))
    g = torch.Generator().manual_seed(21)
    img = torch.randint(0, 256, (3, 150, 221), dtype=torch.uint8, generator=g)
    want = torch.zeros_like(img)
    for _, x, y, win in rs.sliding_window(img, 64):
        o = model(win.unsqueeze(0).contiguous().cuda())[0].cpu()
        want[:, y:y + o.shape[-2], x:x + o.shape[-1]] = o
    assert torch.equal(rs.upscale_image(m, img, 64, "reference"), want)
    assert torch.equal(rs.upscale_image(m, img, 64, "reference", batch=3), want)
    whole = model(img.unsqueeze(0).cuda())[0].cpu()
    assert torch.equal(rs.upscale_image(m, img, 64, "whole"), whole)
    assert torch.equal(rs.upscale_image(m, img, 64, "halo", tile=(64, 96)), whole)
    assert torch.equal(rs.upscale_image(m, img, 64, "halo"), whole)
    small = img[:, :40, :50].contiguous()                       # smaller than the window: one window = the image (rs.py:20-21)
    assert torch.equal(rs.upscale_image(m, small, 64, "reference"), model(small.unsqueeze(0).cuda())[0].cpu())


def test_stream_results_stay_valid_for_depth_more_items(model):
    """a yielded pinned tensor is valid until `depth` more items have been taken (ring of 2 * depth + 1 slots):
    holding depth + 1 results without cloning is safe."""
    g = torch.Generator().manual_seed(17)
    batches = [torch.randint(0, 256, (2, 3, 24, 40), dtype=torch.uint8, generator=g) for _ in range(9)]
    want = [model(b.cuda()).cpu() for b in batches]
    depth = 2
    held = []
    for i, y in enumerate(model.stream(iter(batches), depth=depth)):
        held.append((i, y))                                   # no clone
        for j, t in held[-(depth + 1):]:
            assert torch.equal(t, want[j]), (i, j)
    assert len(held) == len(batches)
