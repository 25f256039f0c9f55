"""CPU tests: the oracle restatement against golden vectors produced by the reference itself
(tools/make_golden.py ran torch.jit.load('/root/reference/model.pt') in the build container)."""
import os

import numpy as np
import pytest
import torch

from oracle import modelpt_oracle as mo
from oracle import l1_oracle


@pytest.fixture(scope="module")
def sd():
    return mo.load_weights()


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "model_pt_golden.npz"))


CASES = ["uniform_2x96x96", "uniform_1x40x56", "uniform_3x17x23", "uniform_1x1x1",
         "smooth_1x72x88", "zeros_1x32x32", "ones255_1x24x40"]


@pytest.mark.parametrize("case", CASES)
def test_oracle_matches_reference_bit_exact(case, sd, gold):
    x = torch.from_numpy(gold[case + ".in"])
    y = mo.forward_u8(x, sd)
    assert y.dtype == torch.uint8
    assert np.array_equal(y.numpy(), gold[case + ".out"])  # 0 LSB


def test_oracle_tanh_prerounding(sd, gold):
    x = torch.from_numpy(gold["uniform_1x40x56.in"])
    t = mo.forward_tanh(x, sd).numpy()
    assert np.max(np.abs(t - gold["uniform_1x40x56.tanh"])) <= 1e-6


def test_weights_inventory(sd):
    live = sum(v.numel() for v in sd.values())
    assert live == 363787
    assert sd["net.1.conv0.0.conv.weight"].shape == (64, 3, 9, 9)
    assert sd["net.1.conv2.0.conv.weight"].shape == (3, 64, 9, 9)
    for i in range(8):
        assert sd[f"net.1.residual.{i}.m.0.conv.weight"].shape == (32, 64, 3, 3)
        assert sd[f"net.1.residual.{i}.m.1.conv.weight"].shape == (64, 32, 3, 3)


def test_rs_tiling_matches_reference(sd, golden_dir):
    g = np.load(os.path.join(golden_dir, "rs_tiling_golden.npz"))
    img = torch.from_numpy(g["image"])
    out = mo.rs_image_reference(img, 96, sd)
    assert np.array_equal(out.numpy(), g["rs_window96"])
    # the seams are real: reference tiling differs from the whole-image forward
    assert not np.array_equal(g["rs_window96"], g["whole"])


def test_halo_tiling_equals_whole_image(sd, golden_dir):
    g = np.load(os.path.join(golden_dir, "rs_tiling_golden.npz"))
    img = torch.from_numpy(g["image"])
    out = mo.halo_tiled(img, 64, 96, mo.RF_RADIUS, sd)
    assert np.array_equal(out.numpy(), g["whole"])  # exact, SURVEY F6


def test_receptive_field_radius(sd):
    g = torch.Generator().manual_seed(7)
    x = torch.randint(0, 256, (1, 3, 80, 80), dtype=torch.uint8, generator=g)
    x2 = x.clone()
    x2[0, :, 40, 40] = 255 - x2[0, :, 40, 40]
    d = (mo.forward_tanh(x, sd) - mo.forward_tanh(x2, sd)).abs().amax(dim=(0, 1))
    ys, xs = torch.nonzero(d > 0, as_tuple=True)
    assert ys.min() >= 40 - mo.RF_RADIUS and ys.max() <= 40 + mo.RF_RADIUS
    assert xs.min() >= 40 - mo.RF_RADIUS and xs.max() <= 40 + mo.RF_RADIUS
    assert ys.max() - ys.min() >= 2 * mo.RF_RADIUS - 6  # and it really is about that wide


def test_l1_oracle_matches_reference(golden_dir):
    g = np.load(os.path.join(golden_dir, "l1_golden.npz"))
    for name in ("a", "b"):
        p, t = g[name + ".pred"], g[name + ".tgt"]
        assert abs(float(l1_oracle.l1_loss(p, t)) - float(g[name + ".loss"])) <= 1e-6 * abs(float(g[name + ".loss"])) + 1e-9
        assert np.array_equal(l1_oracle.l1_grad(p, t), g[name + ".grad"])
    gb = l1_oracle.l1_grad(g["b.pred"], g["b.tgt"])
    assert np.all(gb[0, 0, 0, :3] == 0)  # sign(0) = 0 on exact ties


def test_mse_oracle_matches_torch_mseloss():
    """nn.MSELoss() is the reference's loss without --enchant (train.py:264): value and gradient of the restatement against torch."""
    import torch
    from oracle import l1_oracle
    g = torch.Generator().manual_seed(100)
    p = (torch.rand(4, 3, 32, 48, generator=g) * 2 - 1).requires_grad_(True)
    t = torch.rand(4, 3, 32, 48, generator=g) * 2 - 1
    loss = torch.nn.MSELoss()(p, t)
    (loss * 1024.0).backward()
    assert abs(float(l1_oracle.mse_loss(p.detach().numpy(), t.numpy())) - float(loss)) <= 1e-6 * float(loss)
    got = l1_oracle.mse_grad(p.detach().numpy(), t.numpy(), 1.0, 1024.0)
    assert np.abs(got - p.grad.numpy()).max() <= 1e-6 * np.abs(p.grad.numpy()).max()
