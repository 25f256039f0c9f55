"""Class nets: CPU tests of the oracle against the reference's own outputs and of the drop-in classes'
state-dict compatibility; GPU parity of the drop-in forward."""
import os

import numpy as np
import pytest
import torch

from oracle import nets_oracle as no

CASES = [("eresnet_2_x2", "EResNet", 2, 2, 0.01), ("eresnet_1_x4", "EResNet", 1, 4, 0.01), ("resnet_1_x2", "ResNet", 1, 2, 0.2)]


def _build(cls_name, nb, scale, gold=None, name=None):
    from image_super_resolution_b200 import models
    torch.manual_seed(100)   # the reference's default seed (train.py:160): same construction order => same weights
    net = getattr(models, cls_name)(nb, 0.2, scale).eval()
    if cls_name == "ResNet":  # the golden run randomised the BatchNorm statistics with this generator
        g = torch.Generator().manual_seed(107)
        with torch.no_grad():
            for m in net.modules():
                if isinstance(m, torch.nn.BatchNorm2d):
                    m.running_mean.copy_(torch.randn(m.num_features, generator=g) * 0.1)
                    m.running_var.copy_(torch.rand(m.num_features, generator=g) + 0.5)
                    m.weight.copy_(torch.rand(m.num_features, generator=g) + 0.5)
                    m.bias.copy_(torch.randn(m.num_features, generator=g) * 0.1)
    return net


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "nets_golden.npz"))


@pytest.mark.parametrize("name,cls,nb,scale,s0", CASES)
def test_dropin_state_dict_matches_reference(gold, name, cls, nb, scale, s0):
    net = _build(cls, nb, scale)
    sd = net.state_dict()
    assert list(sd.keys()) == list(gold[name + ".keys"])                     # key-for-key, in order
    sums = np.array([float(v.double().sum()) for v in sd.values()])
    assert np.allclose(sums, gold[name + ".sums"], rtol=0, atol=1e-9)        # same seed => same initial weights
    abss = np.array([float(v.double().abs().sum()) for v in sd.values()])
    assert np.allclose(abss, gold[name + ".abssums"], rtol=0, atol=1e-9)


@pytest.mark.parametrize("name,cls,nb,scale,s0", CASES)
def test_oracle_matches_reference_classes(gold, name, cls, nb, scale, s0):
    net = _build(cls, nb, scale)
    y = no.resnet_forward(torch.from_numpy(gold[name + ".lr"]), net.state_dict(), nb, 0.2, scale, conv0_slope=s0)
    assert np.abs(y.numpy() - gold[name + ".out"]).max() <= 2e-6


@pytest.fixture(scope="module")
def gold16(golden_dir):
    return np.load(os.path.join(golden_dir, "eresnet16_golden.npz"))


def _build16():
    from image_super_resolution_b200 import models
    torch.manual_seed(100)
    return models.EResNet(16, 0.2, 2).eval()


def test_eresnet16_dropin_and_oracle_match_reference(gold16):
    """BASELINE configs[4]'s net, EResNet(16, 0.2, 2) (utils/models.py:621-650): same keys / initial weights as the reference
    class, and the oracle restatement reproduces the reference's forward on [2,3,64,64] (244 convs)."""
    net = _build16()
    sd = net.state_dict()
    assert list(sd.keys()) == list(gold16["keys"]) and len(sd) == 488
    sums = np.array([float(v.double().sum()) for v in sd.values()])
    assert np.allclose(sums, gold16["sums"], rtol=0, atol=1e-9)
    assert sum(p.numel() for p in net.parameters()) == 11726595
    y = no.resnet_forward(torch.from_numpy(gold16["lr"]), sd, 16, 0.2, 2, conv0_slope=0.01)
    assert np.abs(y.numpy() - gold16["out"]).max() <= 2e-6


@pytest.mark.gpu
def test_gpu_eresnet16_forward_matches_reference_golden(gold16):
    """The depth every reported class-net number is quoted on: fp16 operands through 244 convs vs the reference class (fp32).
    Tolerance: SURVEY §8(c) max-abs <= 4e-3 on [0,1] (8e-3 on the tanh output) — and, because EResNet's 0.2-scaled init keeps
    |y| <= 0.022 here, a relative bound of 2 % of max|ref| so that the check is not vacuous."""
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    net = _build16().to("cuda:0")
    with torch.no_grad():
        y = net(torch.from_numpy(gold16["lr"]).to("cuda:0")).cpu().numpy()
    ref = gold16["out"]
    err = np.abs(y - ref).max()
    rel = err / np.abs(ref).max()
    cos = float((y.ravel().astype(np.float64) @ ref.ravel().astype(np.float64)) /
                (np.linalg.norm(y.ravel().astype(np.float64)) * np.linalg.norm(ref.ravel().astype(np.float64))))
    print(f"EResNet(16) x2 forward [2,3,64,64]: max err {err:.3e} ({rel:.3%} of max|ref|), cosine {cos:.6f}")
    assert err <= 8e-3 and rel <= 2e-2, (err, rel)


def test_cpu_inputs_and_training_are_refused():
    from image_super_resolution_b200 import models
    net = models.EResNet(1, 0.2, 2)
    with pytest.raises(RuntimeError):
        net(torch.zeros(1, 3, 8, 8))


@pytest.mark.gpu
@pytest.mark.parametrize("name,cls,nb,scale,s0", CASES)
def test_gpu_forward_matches_reference_golden(gold, name, cls, nb, scale, s0):
    """Tolerance: max-abs <= 4e-3 on the tanh output rescaled to [0,1] (i.e. 8e-3 on [-1,1])."""
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    net = _build(cls, nb, scale).to("cuda:0")
    with torch.no_grad():
        y = net(torch.from_numpy(gold[name + ".lr"]).to("cuda:0")).cpu().numpy()
    ref = gold[name + ".out"]
    assert y.shape == ref.shape
    err = np.abs(y - ref).max()
    assert err <= 8e-3, f"{name}: max err {err}"
    # relative check too: EResNet's 0.2-scaled init makes outputs tiny, an absolute bound alone would be vacuous
    rel = np.abs(y - ref).max() / (np.abs(ref).max() + 1e-12)
    assert rel <= 2e-2, f"{name}: rel err {rel}"


@pytest.mark.gpu
def test_gpu_model_wrapper_uint8_fuse(gold):
    """Model(net).init_normalize + fuse(): uint8 in -> uint8 out equals Normalize -> net -> TanhToArrayImage."""
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    from image_super_resolution_b200 import models
    net = _build("ResNet", 1, 2)
    sd = {k: v.clone() for k, v in net.state_dict().items()}
    m = models.Model(net)
    m.init_normalize([0.485, 0.456, 0.406], [0.229, 0.224, 0.225])
    assert list(m.state_dict().keys())[:3] == ["net.0.mean", "net.0.std", "net.0.max_pixel_value"]
    m.fuse()
    m = m.to("cuda:0").eval()
    g = torch.Generator().manual_seed(1)
    x = torch.randint(0, 256, (2, 3, 20, 28), dtype=torch.uint8, generator=g)
    y = m(x.to("cuda:0")).cpu()
    xn = ((x.float() / 255.0) - torch.tensor([0.485, 0.456, 0.406]).view(1, 3, 1, 1)) / torch.tensor([0.229, 0.224, 0.225]).view(1, 3, 1, 1)
    t = no.resnet_forward(xn, sd, 1, 0.2, 2, conv0_slope=0.2)
    ref = ((t + 1) / 2 * 255).round().to(torch.uint8)
    assert y.dtype == torch.uint8 and y.shape == ref.shape
    assert (y.int() - ref.int()).abs().max().item() <= 2


@pytest.mark.gpu
def test_gpu_captured_forward_replays_correctly():
    """models._NetPlan captures the forward's launches per input shape (CUDA graph with programmatic edges): replays on new
    inputs, other shapes in between and a parameter change (re-pack => re-capture) give what eager launches give, bit for bit."""
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    from image_super_resolution_b200 import models
    dev = torch.device("cuda:0")
    torch.manual_seed(100)
    net = models.EResNet(1, 0.2, 2).to(dev).eval()
    g = torch.Generator().manual_seed(5)
    xs = [torch.randn(2, 3, 24, 40, generator=g).to(dev), torch.randn(2, 3, 24, 40, generator=g).to(dev),
          torch.randn(1, 3, 33, 17, generator=g).to(dev)]
    plan = net._plan(dev)

    def eager(x):
        plan.use_graphs = False
        try:
            return net(x)
        finally:
            plan.use_graphs = True

    with torch.no_grad():
        want = [eager(x) for x in xs]
        for i in (0, 1, 2, 0, 2, 1):
            assert torch.equal(net(xs[i]), want[i]), i
        assert len(plan.graphs) == 2
        net.conv1.conv.weight.mul_(1.5)                      # parameters changed: packed filters and graphs are re-derived
        w2 = eager(xs[0])
        assert not torch.equal(w2, want[0])
        assert torch.equal(net(xs[0]), w2)


# --------------------------------------------------------------------------------------------------------------------------
# the current-revision Denoise generator (utils/models.py:672-706), inference
# --------------------------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def gold_dn(golden_dir):
    return np.load(os.path.join(golden_dir, "denoise_golden.npz"))


def _build_denoise():
    from image_super_resolution_b200 import models
    torch.manual_seed(100)
    net = models.Denoise(4).eval()
    g = torch.Generator().manual_seed(109)       # tools/make_golden.py golden_denoise(): non-trivial BatchNorm statistics
    with torch.no_grad():
        for m in net.modules():
            if isinstance(m, torch.nn.BatchNorm2d):
                m.running_mean.copy_(torch.randn(m.num_features, generator=g) * 0.1)
                m.running_var.copy_(torch.rand(m.num_features, generator=g) + 0.5)
                m.weight.copy_(torch.rand(m.num_features, generator=g) * 0.5 + 0.25)
                m.bias.copy_(torch.randn(m.num_features, generator=g) * 0.1)
    return net


def test_denoise_dropin_and_oracle_match_reference(gold_dn):
    """Same module tree / construction order as the reference class (keys in order, same seeded weights), and the oracle
    restatement (stride-2 conv, 256-channel blocks, bare PixelShuffle + LeakyReLU) reproduces the reference's forward."""
    net = _build_denoise()
    sd = net.state_dict()
    assert list(sd.keys()) == list(gold_dn["keys"])
    sums = np.array([float(v.double().sum()) for v in sd.values()])
    assert np.allclose(sums, gold_dn["sums"], rtol=0, atol=1e-9)
    y = no.denoise_forward(torch.from_numpy(gold_dn["x"]), sd, 4)
    assert np.abs(y.numpy() - gold_dn["out"]).max() <= 2e-6
    assert all(m.inplace for m in net.modules() if hasattr(m, "inplace"))        # utils/models.py:691-693


def test_denoise_refuses_what_is_not_on_the_path():
    from image_super_resolution_b200 import models
    net = models.Denoise(2)
    with pytest.raises(RuntimeError):
        net(torch.zeros(1, 3, 8, 8))                       # CPU tensor
    with pytest.raises(NotImplementedError):
        models.ConvWithoutBN(64, 64, 9, 2)                 # stride 2 only as the 3x3 down-sampling conv


@pytest.mark.gpu
def test_gpu_denoise_forward_matches_reference_golden(gold_dn):
    """fp16 operands / fp32 accumulate through 1 + 4 + 1 + 4 + 4 + 1 + 1 convs vs the reference class (fp32 CPU).
    Tolerance: max-abs <= 8e-3 on the tanh output; the relative bound keeps it from being vacuous."""
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    net = _build_denoise().to("cuda:0")
    x = torch.from_numpy(gold_dn["x"]).to("cuda:0")
    with torch.no_grad():
        y = net(x).cpu().numpy()
        y2 = net(x).cpu().numpy()                          # second call: the captured graph replays
    ref = gold_dn["out"]
    assert y.shape == ref.shape
    err = np.abs(y - ref).max()
    rel = err / np.abs(ref).max()
    print(f"Denoise(4) forward [2,3,32,40]: max err {err:.3e} ({rel:.3%} of max|ref|)")
    assert err <= 8e-3 and rel <= 2e-2, (err, rel)
    assert np.array_equal(y, y2)
    with pytest.raises(NotImplementedError):               # odd sizes / training are not on the sm_100a path
        with torch.no_grad():
            net(torch.zeros(1, 3, 31, 40, device="cuda:0"))
    with pytest.raises(NotImplementedError):
        net(x)                                             # grad enabled, parameters require grad


@pytest.mark.gpu
def test_gpu_denoise_model_wrapper_uint8_fuse_and_larger_frame(gold_dn):
    """Model(Denoise).init_normalize().fuse(): uint8 in -> uint8 out against the oracle on a 96 x 136 frame (several M-tiles per
    row at both resolutions)."""
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    from image_super_resolution_b200 import models
    net = _build_denoise()
    sd = {k: v.clone() for k, v in net.state_dict().items()}
    m = models.Model(net)
    m.init_normalize([0.485, 0.456, 0.406], [0.229, 0.224, 0.225])
    m.fuse()
    m = m.to("cuda:0").eval()
    g = torch.Generator().manual_seed(2)
    x = torch.randint(0, 256, (3, 3, 96, 136), dtype=torch.uint8, generator=g)
    y = m(x.to("cuda:0")).cpu()
    xn = ((x.float() / 255.0) - torch.tensor([0.485, 0.456, 0.406]).view(1, 3, 1, 1)) / torch.tensor([0.229, 0.224, 0.225]).view(1, 3, 1, 1)
    ref = ((no.denoise_forward(xn, sd, 4) + 1) / 2 * 255).round().to(torch.uint8)
    assert y.dtype == torch.uint8 and y.shape == ref.shape
    d = (y.int() - ref.int()).abs()
    assert int(d.max()) <= 2 and float((d == 0).float().mean()) >= 0.97, (int(d.max()), float((d == 0).float().mean()))


# --------------------------------------------------------------------------------------------------------------------------
# ResNet in training mode (BatchNorm batch statistics): utils/models.py:592-618 under train.py:55-57
# --------------------------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def gold_rt(golden_dir):
    return np.load(os.path.join(golden_dir, "resnet_train_golden.npz"))


def _build_resnet_train():
    from image_super_resolution_b200 import models
    torch.manual_seed(100)
    net = models.ResNet(1, 0.2, 2).train()
    g = torch.Generator().manual_seed(111)        # tools/make_golden.py golden_resnet_train()
    with torch.no_grad():
        for m in net.modules():
            if isinstance(m, torch.nn.BatchNorm2d):
                m.weight.copy_(torch.rand(m.num_features, generator=g) + 0.5)
                m.bias.copy_(torch.randn(m.num_features, generator=g) * 0.1)
    return net


def _grad_checks(named_grads, gold_rt, rtol):
    names = list(gold_rt["grad.names"])
    assert [n for n, _ in named_grads] == names
    for i, (name, g) in enumerate(named_grads):
        r = torch.randn(g.numel(), generator=torch.Generator().manual_seed(i), dtype=torch.float64)
        scale = float(gold_rt["grad.abssums"][i]) + 1e-30
        assert abs(float(g.double().sum()) - float(gold_rt["grad.sums"][i])) <= rtol * scale, name
        assert abs(float(g.double().abs().sum()) - scale) <= rtol * scale, name
        assert abs(float(g.double().flatten() @ r) - float(gold_rt["grad.proj"][i])) <= rtol * scale * 4, name


def test_resnet_train_mode_oracle_matches_reference(gold_rt):
    """The oracle's training-mode restatement (batch-statistics BatchNorm, autograd, running-buffer update) against the reference
    class's own training step: output, loss, every gradient (checksums + a random projection) and every BatchNorm buffer."""
    net = _build_resnet_train()
    sd = {k: (v.detach().clone().requires_grad_(True) if v.dtype.is_floating_point and "running" not in k else v.detach().clone())
          for k, v in net.state_dict().items()}
    lr, hr = torch.from_numpy(gold_rt["lr"]), torch.from_numpy(gold_rt["hr"])
    out = no.resnet_forward_train(lr, sd, 1, 0.2, 2, conv0_slope=0.2)
    loss = (out - hr).abs().mean()
    loss.backward()
    assert np.abs(out.detach().numpy() - gold_rt["out"]).max() <= 2e-5
    assert abs(loss.item() - float(gold_rt["loss"])) <= 1e-6
    _grad_checks([(n, sd[n].grad) for n, _ in net.named_parameters()], gold_rt, 2e-4)
    bn = np.concatenate([sd[k].detach().double().numpy().reshape(-1) for k in gold_rt["bn.keys"]])
    assert np.abs(bn - gold_rt["bn.values"]).max() <= 1e-5


@pytest.mark.gpu
def test_gpu_resnet_train_step_matches_reference_and_fp32_autograd(gold_rt):
    """The drop-in ResNet in train() mode on the sm_100a path (training.BNTrainPlan): output and loss against the reference's
    training-mode golden, every gradient against fp32 autograd of the oracle (cosine >= 0.999), the BatchNorm
    running statistics / num_batches_tracked against the reference's — after ONE step although the plan warms up, captures
    and replays the forward — and eval() afterwards uses the moved statistics."""
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    import image_super_resolution_b200 as srb
    dev = torch.device("cuda:0")
    net = _build_resnet_train()
    sd = {k: (v.detach().clone().requires_grad_(True) if v.dtype.is_floating_point and "running" not in k else v.detach().clone())
          for k, v in net.state_dict().items()}
    lr, hr = torch.from_numpy(gold_rt["lr"]), torch.from_numpy(gold_rt["hr"])
    ref_out = no.resnet_forward_train(lr, sd, 1, 0.2, 2, conv0_slope=0.2)
    (ref_out - hr).abs().mean().backward()
    net = net.to(dev).train()
    out = net(lr.to(dev))
    loss = srb.L1Loss()(out, hr.to(dev))
    (loss * 65536.0).backward()
    torch.cuda.synchronize()
    srb._lib.Context.get(0).check_device()
    assert np.abs(out.detach().cpu().numpy() - gold_rt["out"]).max() <= 2e-2
    assert abs(loss.item() - float(gold_rt["loss"])) <= 1e-2 * float(gold_rt["loss"])
    rows = []
    for name, p in net.named_parameters():
        got, ref = p.grad.detach().cpu().double().flatten() / 65536.0, sd[name].grad.double().flatten()
        cos = float(got @ ref / (got.norm() * ref.norm() + 1e-300))
        rows.append((cos, abs(float(got.norm() / (ref.norm() + 1e-300)) - 1), name))
    print(f"ResNet(1) train step: {len(rows)} tensors, gradient cosine min {min(r[0] for r in rows):.5f}, |norm ratio - 1| max "
          f"{max(r[1] for r in rows):.4f}; worst: " + ", ".join(f"{n} {c:.5f}" for c, _, n in sorted(rows)[:4]))
    # conv filters: the 0.999 of the EResNet contract.  BatchNorm gamma / beta gradients are 32- or 64-element vectors, each element
    # a signed sum over all pixels of 16-bit data gradients: measured 0.9987 .. 0.99999 (fp16 operands, the reference's own
    # autocast precision) -> 0.998 for those
    bad = [(n, round(c, 5), round(d, 4)) for c, d, n in rows if c < (0.998 if ".bn." in n else 0.999) or d > 2e-2]
    assert not bad, bad[:10]
    got_sd = net.state_dict()
    for k in gold_rt["bn.keys"]:
        a, b = got_sd[str(k)].detach().cpu().double().reshape(-1), sd[str(k)].detach().double().reshape(-1)
        assert (a - b).abs().max().item() <= 2e-3 * max(1.0, b.abs().max().item()), k
    assert int(net.conv1.bn.num_batches_tracked.item()) == 1          # warm-up + capture + replay = ONE update
    # a second step moves the statistics once more (graph replay), and eval() runs on them
    for p in net.parameters():
        p.grad = None
    out2 = net(lr.to(dev))
    (srb.L1Loss()(out2, hr.to(dev)) * 65536.0).backward()
    assert int(net.conv1.bn.num_batches_tracked.item()) == 2
    ref2 = no.resnet_forward_train(lr, {k: v.detach() for k, v in sd.items()}, 1, 0.2, 2, conv0_slope=0.2)
    assert (out2.detach().cpu() - ref2).abs().max().item() <= 2e-2
    net.eval()
    with torch.no_grad():
        ye = net(lr.to(dev)).cpu()
    ref_e = no.resnet_forward(lr, {k: v.detach() for k, v in sd.items()}, 1, 0.2, 2, conv0_slope=0.2)
    assert (ye - ref_e).abs().max().item() <= 2e-2
