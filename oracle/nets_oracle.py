"""ORACLE (test infrastructure) — CPU fp32 functional restatement of the class-defined SR nets.

Restates, from a plain state_dict, the forward of the reference's `EResNet` / `ResNet`
(utils/models.py:592-650): conv0 9x9 + LeakyReLU -> num_block x RRDB (3 x RDB, utils/models.py:245-271,298-317)
-> conv1 3x3 + long skip -> (scale//2) x Scaler (conv 64->256, PixelShuffle(2), LeakyReLU(0.01); :572-589)
-> conv2 9x9 + tanh.  `ResNet` differs by Conv (bias-free conv + BatchNorm, eval mode = running statistics,
:75-101) inside the RDBs / conv1 and by LeakyReLU(0.2) after conv0 (:596 vs :625).
Pinned by tests/golden/nets_golden.npz: outputs of the reference classes themselves (tools/make_golden.py).
"""
import torch
import torch.nn.functional as F


_BN_TRAIN = [False]     # resnet_forward_train: BatchNorm with batch statistics (nn.BatchNorm2d in training mode)


def _conv(x, sd, prefix, pad):
    w = sd[prefix + ".conv.weight"]
    b = sd.get(prefix + ".conv.bias")
    y = F.conv2d(x, w, b, padding=pad)
    if prefix + ".bn.weight" in sd:  # Conv with BatchNorm (utils/models.py:97-98): eval mode = running statistics;
        # training mode = batch statistics, and the running buffers move by momentum 0.1 (F.batch_norm updates them in place)
        y = F.batch_norm(y, sd[prefix + ".bn.running_mean"], sd[prefix + ".bn.running_var"], sd[prefix + ".bn.weight"],
                         sd[prefix + ".bn.bias"], _BN_TRAIN[0], 0.1 if _BN_TRAIN[0] else 0.0, 1e-5)
        if _BN_TRAIN[0] and prefix + ".bn.num_batches_tracked" in sd:
            sd[prefix + ".bn.num_batches_tracked"] += 1
    return y


def rdb(x, sd, prefix, add_rate, slope=0.01):
    """RDB.forward, utils/models.py:265-271."""
    feats = [x]
    for i in range(4):
        feats.append(F.leaky_relu(_conv(torch.cat(feats, 1), sd, f"{prefix}.conv{i}", 1), slope))
    return _conv(torch.cat(feats, 1), sd, f"{prefix}.conv", 1) * add_rate + x


def rrdb(x, sd, prefix, add_rate):
    """RRDB.forward, utils/models.py:316-317."""
    y = x
    for d in range(3):
        y = rdb(y, sd, f"{prefix}.net.{d}", add_rate)
    return y * add_rate + x


@torch.no_grad()
def resnet_forward(x, sd, num_block, add_rate=0.2, scale=2, conv0_slope=0.01):
    """EResNet.forward (utils/models.py:645-650; conv0_slope=0.01) / ResNet.forward (:613-618; conv0_slope=0.2)."""
    x0 = F.leaky_relu(_conv(x, sd, "conv0", 4), conv0_slope)
    r = x0
    for i in range(num_block):
        r = rrdb(r, sd, f"residual.{i}", add_rate)
    y = x0 + _conv(r, sd, "conv1", 1)
    for k in range(scale // 2):
        y = F.leaky_relu(F.pixel_shuffle(_conv(y, sd, f"scaler.{k}.net.0", 1), 2), 0.01)
    return torch.tanh(_conv(y, sd, "conv2", 4))


def resnet_forward_train(x, sd, num_block, add_rate=0.2, scale=2, conv0_slope=0.2):
    """ResNet.forward (utils/models.py:613-618) in TRAINING mode, differentiable (train.py:55-57): BatchNorm normalises with the
    batch's own mean / biased variance and moves sd's running_mean / running_var / num_batches_tracked in place."""
    _BN_TRAIN[0] = True
    try:
        return resnet_forward.__wrapped__(x, sd, num_block, add_rate, scale, conv0_slope)
    finally:
        _BN_TRAIN[0] = False


def make_lr_hr(seed: int, n: int, hr: int, scale: int = 2):
    """Synthetic (hr, lr) pair of SURVEY §8(d) C5: hr uniform[-1,1]; lr = ImageNet-normalised bicubic downsample."""
    g = torch.Generator().manual_seed(seed)
    hr_t = torch.rand(n, 3, hr, hr, generator=g) * 2 - 1
    lr01 = F.interpolate((hr_t + 1) / 2, scale_factor=1.0 / scale, mode="bicubic", align_corners=False).clamp(0, 1)
    mean = torch.tensor([0.485, 0.456, 0.406]).view(1, 3, 1, 1)
    std = torch.tensor([0.229, 0.224, 0.225]).view(1, 3, 1, 1)
    return hr_t, (lr01 - mean) / std


def _res1(x, sd, prefix, slope=0.2):
    """ResidualBlock1.forward (utils/models.py:208-209): x + m(x), m = Conv(act=LeakyReLU), Conv(act=False) — both with BatchNorm."""
    h = F.leaky_relu(_conv(x, sd, f"{prefix}.m.0", 1), slope)
    return x + _conv(h, sd, f"{prefix}.m.1", 1)


@torch.no_grad()
def denoise_forward(x, sd, residual_blocks):
    """Current-revision Denoise.forward (utils/models.py:697-706; ctor :673-695), eval mode (BatchNorm running statistics)."""
    x0 = F.leaky_relu(_conv(x, sd, "conv0.0", 4), 0.2)
    r = x0
    for i in range(residual_blocks // 2):
        r = _res1(r, sd, f"residual_0.{i}")
    r = F.leaky_relu(F.conv2d(r, sd["residual_conv0.conv.weight"], sd["residual_conv0.conv.bias"], stride=2, padding=1), 0.2)
    for i in range(2):
        r = _res1(r, sd, f"residual_1.{i}")
    r = F.leaky_relu(F.pixel_shuffle(r, 2), 0.2)
    for i in range(residual_blocks // 2):
        r = _res1(r, sd, f"residual_2.{i}")
    y = x0 + _conv(r, sd, "conv1", 1)
    return torch.tanh(_conv(y, sd, "conv2.0", 4))
