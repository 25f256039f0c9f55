"""GPU parity of the fused optimizer tail (train.py:58-63): FlatAdam + GradScaler + clip_grad_norm_ +
attached ModelEMA (two launches over flat arenas) against the reference's own objects — torch.amp.GradScaler,
torch.nn.utils.clip_grad_norm_, torch.optim.Adam and the per-tensor EMA formula of utils/models.py:31-40 — on the same
gradients.  fp32 throughout; tolerance 2e-6 relative to the tensor's max (summation order of the norm differs)."""
import math
import os
import sys

import pytest
import torch
from torch import nn

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    return torch.device("cuda:0")


class _Toy(nn.Module):
    def __init__(self):
        super().__init__()
        self.a = nn.Parameter(torch.randn(7, 5, 3, 3))
        self.b = nn.Parameter(torch.randn(7))
        self.c = nn.Parameter(torch.randn(33, 17))
        self.d = nn.Parameter(torch.randn(1))


def _close(a, b, tol=2e-6):
    return (a - b).abs().max().item() <= tol * max(1.0, b.abs().max().item())


def _set_flat_grads(params, grads):
    """p.grad = views of one zero-padded flat arena, the layout training.TrainPlan.backward leaves."""
    n = sum(p.numel() for p in params)
    arena = torch.zeros((n + 3) // 4 * 4, dtype=torch.float32, device=params[0].device)
    o = 0
    for p, g in zip(params, grads):
        v = arena[o:o + p.numel()].view_as(p)
        v.copy_(g)
        p.grad = v
        o += p.numel()
    return arena


@pytest.mark.parametrize("scale,gmag,wd", [(1024.0, 1.0, 0.0), (65536.0, 300.0, 0.0), (8.0, 0.01, 0.1)])
def test_fused_tail_matches_torch_objects(dev, scale, gmag, wd):
    from image_super_resolution_b200.optim import FlatAdam, GradScaler, clip_grad_norm_
    from image_super_resolution_b200.train import ModelEMA
    torch.manual_seed(1)
    ours, ref = _Toy().to(dev), _Toy().to(dev)
    ref.load_state_dict(ours.state_dict())
    po, pr = list(ours.parameters()), list(ref.parameters())
    opt_o = FlatAdam(po, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=wd)
    opt_r = torch.optim.Adam(pr, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=wd)
    sc_o, sc_r = GradScaler("cuda", init_scale=scale), torch.amp.GradScaler("cuda", init_scale=scale)
    ema_o = ModelEMA(ours, tau=10).attach(opt_o)
    ema_r = ModelEMA(ref, tau=10)
    sched_o = torch.optim.lr_scheduler.LinearLR(opt_o, start_factor=1, end_factor=0.1, total_iters=5)
    sched_r = torch.optim.lr_scheduler.LinearLR(opt_r, start_factor=1, end_factor=0.1, total_iters=5)
    dummy = torch.zeros((), device=dev, requires_grad=True)
    for step in range(4):
        sc_o.scale(dummy.sum()); sc_r.scale(dummy.sum())                     # lazily creates the device scale
        gs = [torch.randn_like(p) * gmag * float(sc_r.get_scale()) for p in pr]     # "scaled" gradients
        opt_o.zero_grad(); opt_r.zero_grad()
        _set_flat_grads(po, gs)
        for p, g in zip(pr, gs):
            p.grad = g.clone()
        sc_o.unscale_(opt_o); sc_r.unscale_(opt_r)
        n_o = clip_grad_norm_(ours.parameters(), 10)
        n_r = torch.nn.utils.clip_grad_norm_(ref.parameters(), 10)
        assert abs(n_o.item() - n_r.item()) <= 2e-6 * n_r.item()
        sc_o.step(opt_o); sc_r.step(opt_r)
        sc_o.update(); sc_r.update()
        sched_o.step(); sched_r.step()
        ema_o.update(ours); ema_r.update(ref)
        for a, b in zip(po, pr):
            assert _close(a.data, b.data), step
        for a, b in zip(ema_o.ema.parameters(), ema_r.ema.parameters()):
            assert _close(a.data, b.data), step
        for a, b in zip(po, pr):
            assert _close(opt_o.state[a]["exp_avg"], opt_r.state[b]["exp_avg"])
            assert _close(opt_o.state[a]["exp_avg_sq"], opt_r.state[b]["exp_avg_sq"], 4e-6)
    assert sc_o.get_scale() == sc_r.get_scale()
    assert float(opt_o.state[po[0]]["step"]) == 4.0


def test_nonfinite_gradient_skips_the_step_on_the_device(dev):
    """GradScaler semantics (train.py:60-61): no optimizer step, the scale backs off, the EMA still blends."""
    from image_super_resolution_b200.optim import FlatAdam, GradScaler, clip_grad_norm_
    from image_super_resolution_b200.train import ModelEMA
    torch.manual_seed(2)
    net = _Toy().to(dev)
    ps = list(net.parameters())
    opt = FlatAdam(ps, lr=1e-2)
    sc = GradScaler("cuda", init_scale=256.0)
    ema = ModelEMA(net, tau=1).attach(opt)
    sc.scale(torch.zeros((), device=dev))
    before = [p.detach().clone() for p in ps]
    e_before = [p.detach().clone() for p in ema.ema.parameters()]
    gs = [torch.randn_like(p) for p in ps]
    gs[2][3, 4] = float("inf")
    _set_flat_grads(ps, gs)
    sc.unscale_(opt)
    clip_grad_norm_(net.parameters(), 10)
    sc.step(opt)
    sc.update()
    ema.update(net)
    assert sc.get_scale() == 128.0
    assert float(opt.state[ps[0]]["step"]) == 0.0
    for p, b in zip(ps, before):
        assert torch.equal(p.data, b)
    d = 0.9999 * (1 - math.exp(-1.0))
    for e, eb, p in zip(ema.ema.parameters(), e_before, ps):
        assert _close(e.data, d * eb + (1 - d) * p.data)
    # the next, finite step is taken
    _set_flat_grads(ps, [torch.randn_like(p) for p in ps])
    opt._stats_valid = False
    sc.unscale_(opt); sc.step(opt); sc.update()
    assert float(opt.state[ps[0]]["step"]) == 1.0
    assert not torch.equal(ps[0].data, before[0])


def test_other_optimizers_and_parameters_fall_through_to_torch(dev):
    from image_super_resolution_b200.optim import GradScaler, clip_grad_norm_
    torch.manual_seed(3)
    net = _Toy().to(dev)
    opt = torch.optim.SGD(net.parameters(), lr=0.1)
    sc = GradScaler("cuda", init_scale=4.0)
    loss = sum((p * p).sum() for p in net.parameters())
    sc.scale(loss).backward()
    sc.unscale_(opt)
    n = clip_grad_norm_(net.parameters(), 1.0)
    ref = torch.sqrt(sum((2 * p.detach()).pow(2).sum() for p in net.parameters()))
    assert abs(n.item() - ref.item()) <= 1e-4 * ref.item()
    sc.step(opt); sc.update()


def test_train_loop_with_fused_tail_matches_torch_tail(dev):
    """The drop-in train() (train.py:41-67) on EResNet with FlatAdam / GradScaler / attached EMA == the same loop with
    torch.optim.Adam / torch.amp.GradScaler / per-tensor EMA, from the same initial weights."""
    from image_super_resolution_b200 import models
    from image_super_resolution_b200 import train as T
    from image_super_resolution_b200.optim import FlatAdam, GradScaler
    from image_super_resolution_b200.training import L1Loss
    from oracle import nets_oracle as no
    hr, lr = no.make_lr_hr(11, 4, 32, 2)
    hr, lr = hr.to(dev), lr.to(dev)
    res = []
    for fused in (False, True):
        torch.manual_seed(100)
        net = models.EResNet(1, 0.2, 2).to(dev)
        if fused:
            opt, sc = FlatAdam(net.parameters(), lr=1e-3), GradScaler("cuda")
            ema = T.ModelEMA(net, tau=5).attach(opt)
        else:
            opt, sc = torch.optim.Adam(net.parameters(), lr=1e-3), torch.amp.GradScaler("cuda")
            ema = T.ModelEMA(net, tau=5)
        sched = torch.optim.lr_scheduler.LinearLR(opt, start_factor=1, end_factor=0.1, total_iters=10)
        losses = T.train(net, ema, [(hr, lr)] * 6, L1Loss().to(dev), opt, sc, sched, 0, None)
        res.append((losses, [p.detach().clone() for p in net.parameters()], [p.detach().clone() for p in ema.ema.parameters()]))
    (l0, p0, e0), (l1, p1, e1) = res
    assert l1[-1] < l1[0]
    assert max(abs(a - b) for a, b in zip(l0, l1)) <= 2e-3 * abs(l0[0])
    # Adam's normalised update turns a last-bit difference of a near-zero gradient into +-lr, and the bf16 re-pack of the
    # weights feeds that back: the two runs are compared on the direction and size of the whole update, not per element
    torch.manual_seed(100)
    init = [p.detach().clone() for p in models.EResNet(1, 0.2, 2).to(dev).parameters()]
    for fin0, fin1 in ((p0, p1), (e0, e1)):
        d0 = torch.cat([(a - i).flatten() for a, i in zip(fin0, init)]).double()
        d1 = torch.cat([(a - i).flatten() for a, i in zip(fin1, init)]).double()
        cos = float((d0 @ d1) / (d0.norm() * d1.norm()))
        assert cos >= 0.995, cos
        assert abs(float(d0.norm() / d1.norm()) - 1.0) <= 0.02


def test_flat_adam_state_dict_interchanges_with_torch_adam(dev):
    """train.py:277-300 saves / restores `optimizer.state_dict()`: a torch.optim.Adam checkpoint loads into FlatAdam (the
    moments land in the flat arenas, the views stay views) and the next step matches torch's; FlatAdam's own state_dict
    loads back into torch.optim.Adam."""
    from image_super_resolution_b200.optim import FlatAdam
    torch.manual_seed(4)
    ref, ours = _Toy().to(dev), _Toy().to(dev)
    ours.load_state_dict(ref.state_dict())
    pr, po = list(ref.parameters()), list(ours.parameters())
    opt_r = torch.optim.Adam(pr, lr=1e-3)
    for _ in range(3):
        for p in pr:
            p.grad = torch.randn_like(p)
        opt_r.step()
    ours.load_state_dict(ref.state_dict())
    opt_o = FlatAdam(po, lr=5e-4)
    opt_o.load_state_dict(opt_r.state_dict())
    assert opt_o.param_groups[0]["lr"] == 1e-3
    assert float(opt_o.state[po[0]]["step"]) == 3.0
    assert opt_o.state[po[0]]["exp_avg"].data_ptr() == opt_o.m_arena.data_ptr()          # still a view of the arena
    gs = [torch.randn_like(p) for p in pr]
    for p, g in zip(pr, gs):
        p.grad = g.clone()
    _set_flat_grads(po, gs)
    opt_r.step(); opt_o.step()
    for a, b in zip(po, pr):
        assert _close(a.data, b.data)
    back = torch.optim.Adam([torch.nn.Parameter(p.detach().clone()) for p in po], lr=1.0)
    back.load_state_dict(opt_o.state_dict())
    for q, b in zip(back.param_groups[0]["params"], pr):
        assert _close(back.state[q]["exp_avg"], opt_r.state[b]["exp_avg"])
        assert float(back.state[q]["step"]) == 4.0


def test_norm_queries_do_not_advance_the_step_counter(dev):
    """the Adam step number advances once per optimizer step TAKEN (in the Adam pass), not whenever the gradient
    statistics are computed — logging the norm before unscale_, after step(), or without a step leaves the bias correction
    alone; a skipped (non-finite) step does not count."""
    from image_super_resolution_b200.optim import FlatAdam, GradScaler, clip_grad_norm_
    torch.manual_seed(3)
    ours, ref = _Toy().to(dev), _Toy().to(dev)
    ref.load_state_dict(ours.state_dict())
    po, pr = list(ours.parameters()), list(ref.parameters())
    opt_o = FlatAdam(po, lr=1e-2)
    opt_r = torch.optim.Adam(pr, lr=1e-2)
    for step in range(3):
        gs = [torch.randn_like(p) for p in pr]
        opt_o.zero_grad(); opt_r.zero_grad()
        _set_flat_grads(po, gs)
        for p, g in zip(pr, gs):
            p.grad = g.clone()
        for _ in range(3):                                   # extra norm queries (logging)
            opt_o._stats_valid = False
            opt_o.total_norm()
        clip_grad_norm_(ours.parameters(), 10)
        torch.nn.utils.clip_grad_norm_(ref.parameters(), 10)
        opt_o.step(); opt_r.step()
        opt_o._stats_valid = False
        opt_o.total_norm()                                   # ... and one after the step
        assert float(opt_o.stats[2].item()) == step + 1
    for a, b in zip(po, pr):
        assert _close(a, b)
    # a non-finite gradient: the step is skipped on the device and not counted
    gs = [torch.randn_like(p) for p in pr]
    gs[1][0] = float("inf")
    opt_o.zero_grad()
    _set_flat_grads(po, gs)
    before = [p.detach().clone() for p in po]
    sc = GradScaler("cuda", init_scale=4.0)
    sc.scale(torch.zeros((), device=dev, requires_grad=True).sum())
    sc.unscale_(opt_o); sc.step(opt_o); sc.update()
    assert float(opt_o.stats[2].item()) == 3
    assert all(torch.equal(a, b) for a, b in zip(po, before))
    # torch's state_dict format round-trips, including an optimizer that has not stepped yet (empty state)
    fresh = FlatAdam(list(_Toy().to(dev).parameters()), lr=1e-2)
    fresh.load_state_dict(torch.optim.Adam(list(_Toy().to(dev).parameters()), lr=1e-2).state_dict())
    assert float(fresh.stats[2].item()) == 0 and float(fresh.m_arena.abs().sum().item()) == 0
    for p in fresh.param_groups[0]["params"]:
        assert fresh.state[p]["exp_avg"].data_ptr() >= fresh.m_arena.data_ptr()


def test_eval_forward_after_fused_steps_uses_the_new_weights(dev):
    """optim.FlatAdam updates the parameters (and the attached EMA arena) through raw pointers, which
    autograd's version counters do not see — the inference plan's cache key includes training.PARAM_EPOCH, so an eval forward
    after k more steps runs on the current weights, for the model and for the EMA copy."""
    import copy
    from image_super_resolution_b200 import models
    from image_super_resolution_b200 import train as T
    from image_super_resolution_b200.training import L1Loss
    from oracle import nets_oracle as no
    torch.manual_seed(100)
    net = models.EResNet(1, 0.2, 2).to(dev)
    hr, lr = no.make_lr_hr(11, 2, 32, 2)
    hr, lr = hr.to(dev), lr.to(dev)
    opt = T.FlatAdam(net.parameters(), lr=1e-3)
    scaler = T.GradScaler("cuda")
    ema = T.ModelEMA(net, tau=5)
    ema.ema.to(dev)
    ema.attach(opt)
    loss_fn = L1Loss().to(dev)

    def fresh_eval(m):
        c = copy.deepcopy(m).eval()                  # a deep copy re-plans from the current parameter values
        with torch.no_grad():
            return c(lr)

    for rnd in range(2):
        T.train(net, ema, [(hr, lr)] * 3, loss_fn, opt, scaler, None, rnd, None)
        net.eval()
        with torch.no_grad():
            y, ye = net(lr), ema.ema(lr)
        assert torch.equal(y, fresh_eval(net)), f"round {rnd}: stale packed filters on the trained model"
        assert torch.equal(ye, fresh_eval(ema.ema)), f"round {rnd}: stale packed filters on the EMA model"
    assert not torch.equal(y, ye)


def test_backward_with_retained_grads_is_refused(dev):
    """the gradients are views of one arena that every backward rewrites; a second backward onto retained .grad
    tensors would alias it (2 g instead of g1 + g2) — refused loudly; zero_grad(set_to_none=False) drops them as well."""
    from image_super_resolution_b200 import models
    from image_super_resolution_b200 import train as T
    from image_super_resolution_b200.training import L1Loss
    from oracle import nets_oracle as no
    torch.manual_seed(100)
    net = models.EResNet(1, 0.2, 2).to(dev).train()
    hr, lr = no.make_lr_hr(12, 2, 32, 2)
    hr, lr = hr.to(dev), lr.to(dev)
    loss_fn = L1Loss().to(dev)
    loss_fn(net(lr), hr).backward()
    g1 = [p.grad.detach().clone() for p in net.parameters()]
    with pytest.raises(RuntimeError, match="accumulated"):
        loss_fn(net(lr), hr).backward()
    opt = T.FlatAdam(net.parameters(), lr=1e-3)
    opt.zero_grad(set_to_none=False)
    assert all(p.grad is None for p in net.parameters())
    loss_fn(net(lr), hr).backward()
    # Same step, same gradients.  Bit-reproducible: every data gradient and every 3x3 weight / bias gradient (one CTA owns a
    # unit, fixed summation order).  NOT bit-reproducible: the 9x9 head / tail weight gradients and their bias gradients, whose
    # pixel tiles are split over CTAs that finish with fp32 red.global.add in arrival order (include/srb200.h: srb_conv9_wgrad)
    # — equal up to fp32 rounding of a different summation order.
    atomics = {"conv0.conv.weight", "conv0.conv.bias", "conv2.conv.weight", "conv2.conv.bias"}
    diffs = []
    for (name, p), g in zip(net.named_parameters(), g1):
        if name in atomics:
            assert (p.grad - g).abs().max().item() <= 1e-5 * g.abs().max().item() + 1e-9, name
        elif not torch.equal(p.grad, g):
            diffs.append((name, float((p.grad - g).abs().max()), float(g.abs().max())))
    assert not diffs, diffs[:8]
