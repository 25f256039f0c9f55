#!/usr/bin/env python
"""Generate golden input/output vectors by running the UNMODIFIED reference in this container.

The reference (/root/reference) has no tests or fixtures of its own, so parity is
pinned on outputs of the reference itself: this script imports / loads it by path, runs it on
seeded synthetic inputs on CPU fp32 and writes small fixtures under tests/golden/.
It cannot run on the GPU box (no /root/reference there); the fixtures travel instead.

Stubs: albumentations, albumentations.pytorch(ToTensorV2), termcolor and
torchvision.io.VideoReader are absent from this image and never executed by the hot path
(SURVEY F11); they are stubbed in sys.modules so `utils.models`, `utils.loss`, `rs` import.

usage: python tools/make_golden.py            (writes tests/golden/*.npz)
"""
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")
SEED = 100  # the reference's default seed, train.py:160


def install_stubs():
    alb = types.ModuleType("albumentations")
    albp = types.ModuleType("albumentations.pytorch")
    albp.ToTensorV2 = object
    alb.pytorch = albp
    for name in ("Compose", "RandomCrop", "Resize", "Normalize", "HorizontalFlip", "VerticalFlip",
                 "ImageOnlyTransform", "BasicTransform", "DualTransform"):
        setattr(alb, name, type(name, (), {"__init__": lambda self, *a, **k: None}))
    sys.modules.setdefault("albumentations", alb)
    sys.modules.setdefault("albumentations.pytorch", albp)
    tc = types.ModuleType("termcolor")
    tc.colored = lambda s, *a, **k: s
    sys.modules.setdefault("termcolor", tc)
    import torchvision.io as tvio
    if not hasattr(tvio, "VideoReader"):
        tvio.VideoReader = object
    if REF not in sys.path:
        sys.path.insert(0, REF)


def smooth_noise_u8(g, n, h, w, sigma=0.03):
    """SURVEY §8(d) 'smooth+noise' generator: bicubic-upsampled 1/8-res uniform + N(0,sigma), clamp, x255."""
    lo = torch.rand(n, 3, max(1, h // 8), max(1, w // 8), generator=g)
    img = torch.nn.functional.interpolate(lo, size=(h, w), mode="bicubic", align_corners=False)
    clean = img.clamp(0, 1)
    noisy = (clean + sigma * torch.randn(n, 3, h, w, generator=g)).clamp(0, 1)
    return (clean * 255).round().to(torch.uint8), (noisy * 255).round().to(torch.uint8)


@torch.no_grad()
def golden_model_pt():
    m = torch.jit.load(os.path.join(REF, "model.pt"), "cpu").eval()
    g = torch.Generator().manual_seed(SEED)
    cases = {}
    x = torch.randint(0, 256, (2, 3, 96, 96), dtype=torch.uint8, generator=g)
    cases["uniform_2x96x96"] = x
    x = torch.randint(0, 256, (1, 3, 40, 56), dtype=torch.uint8, generator=g)
    cases["uniform_1x40x56"] = x
    x = torch.randint(0, 256, (3, 3, 17, 23), dtype=torch.uint8, generator=g)  # ragged: not /8, not /16
    cases["uniform_3x17x23"] = x
    x = torch.randint(0, 256, (1, 3, 1, 1), dtype=torch.uint8, generator=g)  # minimum size
    cases["uniform_1x1x1"] = x
    _, noisy = smooth_noise_u8(g, 1, 72, 88)
    cases["smooth_1x72x88"] = noisy
    cases["zeros_1x32x32"] = torch.zeros(1, 3, 32, 32, dtype=torch.uint8)
    cases["ones255_1x24x40"] = torch.full((1, 3, 24, 40), 255, dtype=torch.uint8)
    out = {}
    net = m.net
    for name, x in cases.items():
        y = m(x.clone())
        assert y.dtype == torch.uint8 and y.shape == x.shape
        out[name + ".in"] = x.numpy()
        out[name + ".out"] = y.numpy()
    # pre-round float (tanh) output of the Denoise sub-module for one case
    x = cases["uniform_1x40x56"]
    t = getattr(net, "1")(getattr(net, "0")(x.clone()))
    out["uniform_1x40x56.tanh"] = t.numpy().astype(np.float32)
    np.savez_compressed(os.path.join(OUT, "model_pt_golden.npz"), **out)
    print("model_pt_golden.npz:", {k: v.shape for k, v in out.items()})
    return m


@torch.no_grad()
def golden_rs_tiling(m):
    """rs.py:82-111 still-image path executed with the reference's own sliding_window."""
    import rs as ref_rs
    g = torch.Generator().manual_seed(SEED + 1)
    image = torch.randint(0, 256, (3, 136, 200), dtype=torch.uint8, generator=g)
    result_image = None
    high = width = 0
    step_size = 96
    for step_size, _, _, win in ref_rs.sliding_window(image, step_size):
        w = m(win.unsqueeze(0).clone())
        if result_image is None:
            result_image = torch.zeros_like(image)
            _, image_high, image_width = result_image.size()
        for frame in w:
            c, h, ww = frame.size()
            result_image[..., high:high + h, width:width + ww] = frame
            width += ww
            if width >= image_width:
                high += h
                width = 0
    whole = m(image.unsqueeze(0).clone())[0]
    np.savez_compressed(os.path.join(OUT, "rs_tiling_golden.npz"), image=image.numpy(),
                        rs_window96=result_image.numpy(), whole=whole.numpy())
    diff = (result_image.int() - whole.int()).abs()
    print("rs_tiling_golden.npz: seam max diff vs whole", int(diff.max()), "frac!=0", float((diff > 0).float().mean()))


@torch.no_grad()
def golden_rs_config1(m):
    """BASELINE configs[0]: rs.py:82-111 on one 256 x 256 RGB image, window 96 (9 windows of 96 / 96 / 64 per axis, batch 1),
    executed with the reference's own sliding_window and stitching loop on the reference's model.pt (CPU fp32)."""
    import rs as ref_rs
    g = torch.Generator().manual_seed(SEED)
    image = torch.randint(0, 256, (3, 256, 256), dtype=torch.uint8, generator=g)
    result_image = None
    high = width = 0
    for step_size, _, _, win in ref_rs.sliding_window(image, 96):
        w = m(win.unsqueeze(0).clone())
        if result_image is None:
            result_image = torch.zeros_like(image)
            _, image_high, image_width = result_image.size()
        for frame in w:
            c, h, ww = frame.size()
            result_image[..., high:high + h, width:width + ww] = frame
            width += ww
            if width >= image_width:
                high += h
                width = 0
    whole = m(image.unsqueeze(0).clone())[0]
    np.savez_compressed(os.path.join(OUT, "rs_config1_golden.npz"), image=image.numpy(), rs_window96=result_image.numpy(),
                        whole=whole.numpy())
    print("rs_config1_golden.npz: 256x256, window 96; seam max diff vs whole", int((result_image.int() - whole.int()).abs().max()))


def golden_l1():
    from utils.loss import L1Loss
    g = torch.Generator().manual_seed(SEED + 2)
    out = {}
    crit = L1Loss(1.0)
    for name, shape in (("a", (2, 3, 16, 16)), ("b", (1, 3, 7, 5)), ("c", (4, 3, 128, 128))):
        p = (torch.rand(shape, generator=g) * 2 - 1).requires_grad_(True)
        t = torch.rand(shape, generator=g) * 2 - 1
        if name == "b":
            with torch.no_grad():
                p[0, 0, 0, :3] = t[0, 0, 0, :3]  # exact ties: sign(0) = 0
        loss = crit(p, t)
        loss.backward()
        if name != "c":
            out[name + ".pred"], out[name + ".tgt"] = p.detach().numpy(), t.numpy()
            out[name + ".grad"] = p.grad.numpy()
        else:  # large case: keep only seeds' statistics
            out[name + ".grad_abs_sum"] = np.float64(p.grad.abs().double().sum().item())
        out[name + ".loss"] = np.float32(loss.item())
    np.savez_compressed(os.path.join(OUT, "l1_golden.npz"), **out)
    print("l1_golden.npz:", {k: getattr(v, 'shape', ()) for k, v in out.items()})


@torch.no_grad()
def golden_nets():
    """Reference EResNet / ResNet (utils/models.py) forward on seeded inputs; weights are NOT stored: the drop-in
    classes re-create them from the same seed (same construction order), so only checksums + outputs are kept."""
    from utils.models import EResNet, ResNet
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
    from oracle.nets_oracle import make_lr_hr
    out = {}
    for name, cls, nb, scale in (("eresnet_2_x2", EResNet, 2, 2), ("eresnet_1_x4", EResNet, 1, 4), ("resnet_1_x2", ResNet, 1, 2)):
        torch.manual_seed(SEED)
        net = cls(nb, 0.2, scale).eval()
        if cls is ResNet:  # give BatchNorm non-trivial running statistics and affine parameters
            g = torch.Generator().manual_seed(SEED + 7)
            for m in net.modules():
                if isinstance(m, torch.nn.BatchNorm2d):
                    m.running_mean.copy_(torch.randn(m.num_features, generator=g) * 0.1)
                    m.running_var.copy_(torch.rand(m.num_features, generator=g) + 0.5)
                    m.weight.copy_(torch.rand(m.num_features, generator=g) + 0.5)
                    m.bias.copy_(torch.randn(m.num_features, generator=g) * 0.1)
        _, lr = make_lr_hr(SEED + 3, 2, 24 * scale, scale)   # LR 24x24
        y = net(lr.clone())
        sd = net.state_dict()
        out[name + ".lr"] = lr.numpy()
        out[name + ".out"] = y.numpy()
        out[name + ".keys"] = np.array(list(sd.keys()))
        out[name + ".sums"] = np.array([float(v.double().sum()) for v in sd.values()])
        out[name + ".abssums"] = np.array([float(v.double().abs().sum()) for v in sd.values()])
        print(name, "params", sum(p.numel() for p in net.parameters()), "out", tuple(y.shape), float(y.abs().mean()))
    np.savez_compressed(os.path.join(OUT, "nets_golden.npz"), **out)


@torch.no_grad()
def golden_eresnet16():
    """BASELINE configs[4]'s net: the reference EResNet(16, 0.2, 2) (utils/models.py:621-650; 244 convs) forward on a seeded
    [2,3,64,64] LR batch.  Weights are re-created by the drop-in from the same seed; only checksums travel."""
    from utils.models import EResNet
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
    from oracle.nets_oracle import make_lr_hr
    torch.manual_seed(SEED)
    net = EResNet(16, 0.2, 2).eval()
    _, lr = make_lr_hr(SEED + 5, 2, 128, 2)
    y = net(lr.clone())
    sd = net.state_dict()
    out = {"lr": lr.numpy(), "out": y.numpy(), "keys": np.array(list(sd.keys())),
           "sums": np.array([float(v.double().sum()) for v in sd.values()])}
    np.savez_compressed(os.path.join(OUT, "eresnet16_golden.npz"), **out)
    print("eresnet16_golden.npz: params", sum(p.numel() for p in net.parameters()), "out", tuple(y.shape), "mean|y|", float(y.abs().mean()),
          "max|y|", float(y.abs().max()))


@torch.no_grad()
def golden_denoise():
    """The CURRENT-revision Denoise generator (utils/models.py:672-706: stride-2 3x3 64->256, 256-channel ResidualBlock1s, bare
    PixelShuffle(2) + LeakyReLU) in eval mode with non-trivial BatchNorm statistics, forward on a seeded [2,3,32,40] batch."""
    from utils.models import Denoise
    torch.manual_seed(SEED)
    net = Denoise(4).eval()
    g = torch.Generator().manual_seed(SEED + 9)
    for m in net.modules():
        if isinstance(m, torch.nn.BatchNorm2d):
            m.running_mean.copy_(torch.randn(m.num_features, generator=g) * 0.1)
            m.running_var.copy_(torch.rand(m.num_features, generator=g) + 0.5)
            m.weight.copy_(torch.rand(m.num_features, generator=g) * 0.5 + 0.25)
            m.bias.copy_(torch.randn(m.num_features, generator=g) * 0.1)
    x = torch.randn(2, 3, 32, 40, generator=g)
    y = net(x.clone())
    sd = net.state_dict()
    out = {"x": x.numpy(), "out": y.numpy(), "keys": np.array(list(sd.keys())),
           "sums": np.array([float(v.double().sum()) for v in sd.values()])}
    np.savez_compressed(os.path.join(OUT, "denoise_golden.npz"), **out)
    print("denoise_golden.npz: params", sum(p.numel() for p in net.parameters()), "keys", len(sd), "out", tuple(y.shape), "mean|y|", float(y.abs().mean()))


def golden_resnet_train():
    """One TRAINING step's forward + backward of the reference `ResNet(1, 0.2, 2)` (utils/models.py:592-618) in train mode —
    BatchNorm batch statistics, running-statistics update, autograd through it — with nn.L1Loss on a seeded batch, CPU fp32.
    Stored: the output, the loss, every updated BatchNorm buffer, and per-gradient checksums (sum, |sum|, a seeded random
    projection): the drop-in re-creates the weights from the same seed."""
    from utils.models import ResNet
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
    from oracle.nets_oracle import make_lr_hr
    torch.manual_seed(SEED)
    net = ResNet(1, 0.2, 2).train()
    g = torch.Generator().manual_seed(SEED + 11)
    with torch.no_grad():
        for m in net.modules():
            if isinstance(m, torch.nn.BatchNorm2d):
                m.weight.copy_(torch.rand(m.num_features, generator=g) + 0.5)
                m.bias.copy_(torch.randn(m.num_features, generator=g) * 0.1)
    hr, lr = make_lr_hr(SEED + 13, 4, 48, 2)          # LR 24 x 24, B = 4
    y = net(lr.clone())
    loss = torch.nn.L1Loss()(y, hr)
    loss.backward()
    out = {"lr": lr.numpy(), "hr": hr.numpy(), "out": y.detach().numpy(), "loss": np.float32(loss.item())}
    names, sums, abss, proj = [], [], [], []
    for name, p in net.named_parameters():
        r = torch.randn(p.numel(), generator=torch.Generator().manual_seed(len(names)), dtype=torch.float64)
        names.append(name); sums.append(float(p.grad.double().sum())); abss.append(float(p.grad.double().abs().sum()))
        proj.append(float(p.grad.double().flatten() @ r))
    out["grad.names"], out["grad.sums"], out["grad.abssums"], out["grad.proj"] = np.array(names), np.array(sums), np.array(abss), np.array(proj)
    bn = {k: v.numpy() for k, v in net.state_dict().items() if "running_" in k or "num_batches" in k}
    out["bn.keys"] = np.array(list(bn.keys()))
    out["bn.values"] = np.concatenate([np.asarray(v, dtype=np.float64).reshape(-1) for v in bn.values()])
    np.savez_compressed(os.path.join(OUT, "resnet_train_golden.npz"), **out)
    print("resnet_train_golden.npz: loss", float(loss), "tensors", len(names), "bn buffers", len(bn))


if __name__ == "__main__":
    torch.set_num_threads(os.cpu_count() or 1)
    install_stubs()
    os.makedirs(OUT, exist_ok=True)
    m = golden_model_pt()
    golden_rs_tiling(m)
    golden_rs_config1(m)
    golden_l1()
    golden_nets()
    golden_eresnet16()
    golden_denoise()
    golden_resnet_train()
