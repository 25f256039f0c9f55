"""Drop-in model classes with the reference's names, constructor arguments and state-dict keys
(utils/models.py: Conv :75, ConvWithoutBN :174, RDB :245, RRDB :298, Scaler :572, ResNet :592, EResNet :621,
SRGAN :653, Model :723) whose forward runs on the sm_100a kernels of libsrb200.so.

The sub-modules are parameter containers built in the reference's construction order (so the same
`torch.manual_seed` gives the same initial weights and checkpoints interchange key-for-key); the arithmetic of a
whole net is issued by `_NetPlan` as a fixed schedule of fused conv launches over C8-plane buffers:
  * RDB: the five convs read / write plane slices of ONE 192-channel buffer (torch.cat is free),
  * `*add_rate + x` (RDB) and the RRDB skip are epilogue terms (two residuals on the last conv of an RRDB),
  * Scaler: conv 64->256 as four 64-channel launches with nn.PixelShuffle(2) folded into the store,
  * BatchNorm of `ResNet` in eval mode is folded into the packed filter (fuse_conv_and_bn semantics).
Training (autograd) goes through training.TrainPlan; there is no PyTorch fallback: CPU tensors and train-mode
BatchNorm are refused.
"""
from __future__ import annotations

import torch
from torch import nn

from . import ops, _lib, training as _training
from ._lib import Context


def _act(act):
    """utils/general.py:28-37 without the eval(): a fresh module per use, True -> SiLU, falsy -> Identity."""
    if act is True:
        return nn.SiLU()
    if isinstance(act, nn.LeakyReLU):
        return nn.LeakyReLU(act.negative_slope)
    if isinstance(act, nn.Tanh):
        return nn.Tanh()
    if isinstance(act, nn.Module):
        return act
    return nn.Identity()


class Conv(nn.Module):
    """utils/models.py:75-111 — bias-free conv + BatchNorm + activation (parameter container)."""

    def __init__(self, c1, c2, k=1, s=1, p=None, g=1, d=1, act=True, dropout=0.0):
        super().__init__()
        if s != 1 or g != 1 or d != 1 or dropout != 0.0:
            raise NotImplementedError("hot-path convs are stride 1, groups 1, dilation 1, dropout 0 (SURVEY a8)")
        self.conv = nn.Conv2d(c1, c2, k, s, k // 2 if p is None else p, bias=False)
        self.bn = nn.BatchNorm2d(c2)
        self.drop = nn.Identity()
        self.act = _act(act)
        self.store_bn = nn.Identity()

    def fuseforward(self):
        if isinstance(self.bn, nn.BatchNorm2d):
            self.store_bn = self.bn
            self.bn = nn.Identity()

    def defuseforward(self):
        if isinstance(self.bn, nn.Identity):
            self.bn = self.store_bn
            self.store_bn = nn.Identity()


class ConvWithoutBN(nn.Module):
    """utils/models.py:174-199 — biased conv + activation (parameter container)."""

    def __init__(self, c1, c2, k=1, s=1, p=None, g=1, d=1, act=True, dropout=0.0):
        super().__init__()
        # stride 2 only as Denoise's down-sampling 3x3 conv (utils/models.py:680), run as a stride-1 conv over the
        # space-to-depth input (_DenoisePlan)
        if s not in (1, 2) or (s == 2 and k != 3) or g != 1 or d != 1 or dropout != 0.0:
            raise NotImplementedError("hot-path convs are stride 1 (3x3: or 2), groups 1, dilation 1, dropout 0 (SURVEY a8)")
        self.conv = nn.Conv2d(c1, c2, k, s, k // 2 if p is None else p, bias=True)
        self.drop = nn.Identity()
        self.act = _act(act)


class ResidualBlock1(nn.Module):
    """utils/models.py:202-209 — x + Conv(Conv(x)) (parameter container)."""

    def __init__(self, in_channel, out_channel, hidden_channel, kernel, act):
        super().__init__()
        self.m = nn.Sequential(Conv(in_channel, hidden_channel, kernel, 1, None, act=act),
                               Conv(hidden_channel, out_channel, kernel, 1, None, act=False))


class RDB(nn.Module):
    """utils/models.py:245-271."""

    def __init__(self, in_channel, growth_channel, kernel_size, act, add_rate=0.0, use_BN=True):
        super().__init__()
        self.add_rate = add_rate
        C_ = Conv if use_BN else ConvWithoutBN
        for i in range(4):
            setattr(self, f"conv{i}", C_(in_channel + growth_channel * i, growth_channel, kernel_size, 1, None, act=act))
        self.conv = C_(in_channel + growth_channel * 4, in_channel, kernel_size, 1, None, act=False)


class RRDB(nn.Module):
    """utils/models.py:298-317."""

    def __init__(self, filters, kernel, act, add_rate=0.2, use_BN=True):
        super().__init__()
        assert 0 < add_rate <= 1, "add must be in range (0, 1]"
        self.net = nn.Sequential(*[RDB(filters, filters // 2, kernel, act, add_rate=add_rate, use_BN=use_BN)
                                   for _ in range(3)])
        self.add_rate = add_rate


class Scaler(nn.Module):
    """utils/models.py:572-589."""

    def __init__(self, in_channel, out_channel, scale_factor, kernel_size, act):
        super().__init__()
        if scale_factor != 2:
            raise NotImplementedError("the reference only builds PixelShuffle(2) scalers (utils/models.py:603,632)")
        self.net = nn.Sequential(ConvWithoutBN(in_channel, out_channel * scale_factor ** 2, kernel_size, 1, None, act=False),
                                 nn.PixelShuffle(scale_factor), _act(act))


def _slope(act_module) -> tuple[float, int]:
    """(slope, flag) of the activation following a conv: LeakyReLU -> fused, Identity -> none."""
    if isinstance(act_module, nn.LeakyReLU):
        return float(act_module.negative_slope), ops.F_LRELU
    if isinstance(act_module, (nn.Identity, nn.Tanh)):
        return 0.0, 0
    raise NotImplementedError(f"activation {act_module.__class__.__name__} has no fused epilogue")


class _NetPlan:
    """Packed filters + buffers + launch schedule of one ResNet/EResNet on one device (inference forward)."""

    def __init__(self, net: "_SRNet", device: torch.device, dtype):
        self.device, self.dtype = device, ops.dtype_code(dtype)
        self.ctx = Context.get(device.index)
        self.versions = None
        self.net = net
        self.buffers = {}
        self.use_graphs = True      # the ~260 launches of a forward are captured once per input shape (host-bound otherwise)
        self.graphs = {}
        self.repack()

    def _pack(self, m, cout_slice=None):
        w = m.conv.weight.detach()
        b = m.conv.bias.detach() if m.conv.bias is not None else None
        bn = None
        if isinstance(getattr(m, "bn", None), nn.BatchNorm2d):
            if m.bn.training:
                raise NotImplementedError("train-mode BatchNorm (batch statistics) is not on the sm_100a path yet; "
                                          "call .eval() (SURVEY §8(f) rank 3)")
            bn = (m.bn.weight.detach(), m.bn.bias.detach(), m.bn.running_mean, m.bn.running_var)
            eps = m.bn.eps
        else:
            eps = 1e-5
        if cout_slice is not None:
            w = w[cout_slice]
            b = None if b is None else b[cout_slice]
            bn = None if bn is None else tuple(t[cout_slice] for t in bn)
        return ops.PackedConv(self.ctx, w, b, bn=bn, bn_eps=eps, dtype=self.dtype)

    def _param_versions(self):
        # autograd's version counters do not see writes through raw pointers (optim.FlatAdam's fused step, its attached EMA
        # arena) or through `.data` (broadcast_parameters): those bump training.PARAM_EPOCH instead
        return (tuple(p._version for p in self.net.parameters()) + tuple(b._version for b in self.net.buffers())
                + (_training.PARAM_EPOCH[0],))

    def repack(self):
        n = self.net
        with torch.cuda.device(self.device):
            self.conv0 = self._pack(n.conv0)
            self.rdbs = [[[self._pack(getattr(rdb, f"conv{i}")) for i in range(4)] + [self._pack(rdb.conv)]
                          for rdb in rrdb.net] for rrdb in n.residual]
            self.conv1 = self._pack(n.conv1)
            self.scalers = [[self._pack(s.net[0], slice(64 * q, 64 * q + 64)) for q in range(4)] for s in n.scaler]
            self.conv2 = self._pack(n.conv2)
        self.versions = self._param_versions()
        self.graphs = {}            # captured launches hold the addresses of the previous packed filters

    def _buf(self, name, N, planes, H, W):
        key = (name, N, planes, H, W)
        if key not in self.buffers:
            if not self.graphs:     # (captured graphs of other shapes still use their buffers)
                self.buffers = {k: v for k, v in self.buffers.items() if k[0] != name}
            self.buffers[key] = ops.c8_empty(N, planes, H, W, self.dtype, self.device)
        return self.buffers[key]

    @torch.no_grad()
    def forward(self, x: torch.Tensor | None, u8_out: bool = False, fill_input=None, shape=None) -> torch.Tensor:
        """x: fp32 [N,3,H,W] (converted to the C8 input plane by the library's kernel), or `fill_input(plane)` writes the
        plane itself (Model.forward: Normalize fused into the uint8 -> C8 pack).  Every kernel of the forward is the library's:
        the input conversion writes straight into the captured graph's static input plane, the trunk replays as one graph,
        and the tail conv is launched into a fresh output tensor (no staging copies)."""
        if self.versions != self._param_versions():
            self.repack()   # parameters changed (optimizer step / load_state_dict): derived caches are re-derived
        N, _, H, W = shape if shape is not None else x.shape

        def fill(plane):
            if fill_input is not None:
                fill_input(plane)
            else:
                ops.nchw_to_c8(x, out=plane, planes=1, dtype=self.dtype)

        if not self.use_graphs:
            xin = self._buf("xin", N, 1, H, W)
            fill(xin)
            cur, h, w = self._run_trunk(xin, N, H, W)
            return self._run_tail(cur, N, h, w, u8_out)
        key = (N, H, W)
        st = self.graphs.get(key)
        if st is None:
            if len(self.graphs) >= 8:                     # many different shapes (ragged tiles): start over
                self.graphs, self.buffers = {}, {}
            xs = ops.c8_empty(N, 1, H, W, self.dtype, self.device)
            fill(xs)
            self._run_trunk(xs, N, H, W)                  # eager once: buffers, kernel attributes
            torch.cuda.synchronize(self.device)
            g = torch.cuda.CUDAGraph()
            _lib.check(_lib.lib().srb_set_programmatic_launch(self.ctx.handle, 1), "srb_set_programmatic_launch")
            try:                                          # the filters are constants of the graph: its conv nodes are chained
                with torch.cuda.graph(g, capture_error_mode="thread_local"):     # by programmatic dependent launch edges
                    cur, h, w = self._run_trunk(xs, N, H, W)
            finally:
                _lib.check(_lib.lib().srb_set_programmatic_launch(self.ctx.handle, 0), "srb_set_programmatic_launch")
            st = self.graphs[key] = (g, xs, cur, h, w)
        g, xs, cur, h, w = st
        fill(xs)
        g.replay()
        return self._run_tail(cur, N, h, w, u8_out)

    def _run_trunk(self, xin: torch.Tensor, N: int, H: int, W: int):
        """conv0 .. last Scaler; returns (tail input, h, w)."""
        n = self.net
        # X holds conv0's output (the long skip, utils/models.py:647) in planes 0..7 for the whole forward and doubles as the
        # first RDB's dense buffer; three more dense buffers rotate through the RRDBs
        X = self._buf("denseX", N, 24, H, W)
        rot = [self._buf(f"dense{i}", N, 24, H, W) for i in range(3)]
        s0, f0 = _slope(n.conv0.act)
        ops.conv2d_fwd(self.conv0, xin, 1, 0, X, 24, 0, N, H, W, slope=s0, flags=f0)
        a, free = X, list(rot)                         # a: buffer holding the current RRDB's input (its skip)
        for rrdb, packs in zip(n.residual, self.rdbs):
            b, c = free[0], free[1]                    # RDB1: a->b, RDB2: b->c, RDB3: c->b
            src_dst = [(a, b), (b, c), (c, b)]
            for d, (rdb, pk) in enumerate(zip(rrdb.net, packs)):
                src, dst = src_dst[d]
                for i in range(4):
                    si, fi = _slope(getattr(rdb, f"conv{i}").act)
                    ops.conv2d_fwd(pk[i], src, 24, 0, src, 24, 8 + 4 * i, N, H, W, slope=si, flags=fi)
                r = float(rdb.add_rate)
                if d < 2:     # add_rate * conv + rdb_in  (utils/models.py:271)
                    ops.conv2d_fwd(pk[4], src, 24, 0, dst, 24, 0, N, H, W, residual=src, res_planes=24, res_off=0,
                                   res_scale=1.0, out_scale=r)
                else:         # + the RRDB skip: R*(r*conv + rdb_in) + rrdb_in  (utils/models.py:317)
                    R = float(rrdb.add_rate)
                    ops.conv2d_fwd(pk[4], src, 24, 0, dst, 24, 0, N, H, W, residual=src, res_planes=24, res_off=0,
                                   res_scale=R, out_scale=R * r, res2=a, res2_planes=24, res2_off=0, res2_scale=1.0)
            free = [t for t in rot if t is not b]      # the output (b) is the next RRDB's input; X is never recycled
            a = b
        t = self._buf("t", N, 8, H, W)
        ops.conv2d_fwd(self.conv1, a, 24, 0, t, 8, 0, N, H, W, residual=X, res_planes=24, res_off=0)
        cur, h, w = t, H, W
        for k, (s, pk) in enumerate(zip(n.scaler, self.scalers)):
            up = self._buf(f"up{k}", N, 8, 2 * h, 2 * w)
            sl, fl = _slope(s.net[2])
            for q in range(4):
                ops.conv2d_fwd(pk[q], cur, 8, 0, up, 8, 2 * q, N, h, w, slope=sl, flags=fl | ops.F_PIXSHUF2)
            cur, h, w = up, 2 * h, 2 * w
        return cur, h, w

    def _run_tail(self, cur: torch.Tensor, N: int, h: int, w: int, u8_out: bool) -> torch.Tensor:
        tanh = ops.F_TANH if isinstance(self.net.conv2.act, nn.Tanh) else 0
        if u8_out:
            out = torch.empty((N, 3, h, w), dtype=torch.uint8, device=self.device)
            ops.conv2d_fwd(self.conv2, cur, 8, 0, out, 0, 0, N, h, w, flags=tanh | ops.F_U8OUT)
        else:
            out = torch.empty((N, 3, h, w), dtype=torch.float32, device=self.device)
            ops.conv2d_fwd(self.conv2, cur, 8, 0, out, 0, 0, N, h, w, flags=tanh | ops.F_F32OUT)
        return out

    def launches(self) -> int:
        return 2 + 5 * 3 * len(self.rdbs) + 1 + 4 * len(self.scalers) + 1


class _SRNet(nn.Module):
    """Common forward of ResNet / EResNet (utils/models.py:613-618, :645-650)."""
    operand_dtype = "fp16"
    trainable = True

    plan_class = _NetPlan

    def _plan(self, device) -> _NetPlan:
        plans = self.__dict__.setdefault("_plans", {})
        if device not in plans:
            plans[device] = self.plan_class(self, device, self.operand_dtype)
        return plans[device]

    def forward(self, inputs: torch.Tensor) -> torch.Tensor:
        if not inputs.is_cuda:
            raise RuntimeError("the B200 path has no CPU fallback: move the model and its inputs to a CUDA device")
        if inputs.dim() != 4 or inputs.shape[1] != 3:
            raise RuntimeError(f"expected [N,3,h,w], got {tuple(inputs.shape)}")
        if torch.is_grad_enabled() and self._any_requires_grad() and not self.trainable:
            raise NotImplementedError(f"{self.__class__.__name__}: only the inference forward is on the sm_100a path "
                                      "(wrap the call in torch.no_grad() or .requires_grad_(False); SURVEY §8(f))")
        if torch.is_grad_enabled() and self._any_requires_grad():
            # train.py:55 — forward with saved activations; backward runs on the same kernels (training.py)
            from .training import train_forward
            return train_forward(self, inputs)
        return self._plan(inputs.device).forward(inputs.float().contiguous())

    def _any_requires_grad(self) -> bool:
        # (the module-tree walk of self.parameters() costs 0.15 ms per call: it sits on the host path between two steps)
        ps = self.__dict__.get("_param_list")
        if ps is None or len(ps[1]) != ps[0]:
            lst = list(self.parameters())
            ps = self.__dict__["_param_list"] = (len(lst), lst)
        return any(p.requires_grad for p in ps[1])

    def __deepcopy__(self, memo):  # ModelEMA deep-copies the model (utils/models.py:25): plans are derived caches
        import copy
        cls = self.__class__
        new = cls.__new__(cls)
        memo[id(self)] = new
        for k, v in self.__dict__.items():
            if k not in ("_plans", "_train_plans", "_param_list"):
                new.__dict__[k] = copy.deepcopy(v, memo)
        return new


class ResNet(_SRNet):
    """utils/models.py:592-618."""

    def __init__(self, num_block_resnet=16, add_rate=0.2, scaleRate=2):
        super().__init__()
        scaleRate = scaleRate // 2
        self.conv0 = ConvWithoutBN(3, 64, 9, 1, None, act=nn.LeakyReLU(0.2))
        self.residual = nn.Sequential(*[RRDB(64, 3, act=nn.LeakyReLU(), add_rate=add_rate) for _ in range(num_block_resnet)])
        self.conv1 = Conv(64, 64, 3, 1, None, act=False)
        self.scaler = nn.Sequential(*[Scaler(64, 64, 2, 3, nn.LeakyReLU()) for _ in range(scaleRate)])
        self.conv2 = ConvWithoutBN(64, 3, 9, 1, act=nn.Tanh())


class EResNet(_SRNet):
    """utils/models.py:621-650 (no BatchNorm; every conv weight scaled by 0.2 at construction, :638-640)."""

    def __init__(self, num_block_resnet=16, add_rate=0.2, scaleRate=2):
        super().__init__()
        scaleRate = scaleRate // 2
        self.conv0 = ConvWithoutBN(3, 64, 9, 1, None, act=nn.LeakyReLU())
        self.residual = nn.Sequential(*[RRDB(64, 3, act=nn.LeakyReLU(), add_rate=add_rate, use_BN=False)
                                        for _ in range(num_block_resnet)])
        self.conv1 = ConvWithoutBN(64, 64, 3, 1, None, act=False)
        self.scaler = nn.Sequential(*[Scaler(64, 64, 2, 3, nn.LeakyReLU()) for _ in range(scaleRate)])
        self.conv2 = ConvWithoutBN(64, 3, 9, 1, act=nn.Tanh())
        for x in self.modules():
            if isinstance(x, nn.Conv2d):
                x.weight.data *= 0.2


class _DenoisePlan(_NetPlan):
    """Launch schedule of the current-revision `Denoise` generator (utils/models.py:672-706), inference (BatchNorm folded).

      conv0 9x9 + LReLU                                      head kernel (as the SR nets)
      residual_0 / residual_2: x + Conv(LReLU(Conv(x)))      two 3x3 64->64 launches per block, the skip in the second's epilogue
      residual_conv0: 3x3 stride 2, 64->256, LReLU            space-to-depth (srb_pixunshuffle_mask, slope 1: channel 4c+2i+j of
                                                             pixel (y,x) = channel c of pixel (2y+i, 2x+j)) + a STRIDE-1 3x3 conv
                                                             256->256 at H/2 x W/2 whose filter holds the nine taps at
                                                             (DY, i) = (0,1), (1,0), (1,1) for dy = 0, 1, 2 (same for dx), zeros
                                                             elsewhere: exactly the reference's sums, zero padding included
      residual_1: 256->256 blocks                            weight-streaming kernels, four 64-output launches per conv
      residual_conv1: PixelShuffle(2) + LReLU                srb_pixshuffle_lrelu
      conv1 + long skip, conv2 9x9 + tanh                    as the SR nets
    """

    def _pack_down(self, m, q):
        w = m.conv.weight.detach().float().cpu().numpy()          # [256][64][3][3], stride 2, padding 1
        co, ci = w.shape[:2]
        import numpy as np
        w2 = np.zeros((co, ci, 2, 2, 3, 3), dtype=np.float32)     # [o][c][i][j][DY][DX]
        tap = ((0, 1), (1, 0), (1, 1))                            # d -> (D, parity): input row 2y + d - 1 = 2(y + D - 1) + parity
        for dy, (DY, i) in enumerate(tap):
            for dx, (DX, j) in enumerate(tap):
                w2[:, :, i, j, DY, DX] = w[:, :, dy, dx]
        w2 = w2.reshape(co, ci * 4, 3, 3)
        sl = slice(64 * q, 64 * q + 64)
        return ops.PackedConv(self.ctx, w2[sl], m.conv.bias.detach()[sl], dtype=self.dtype)

    def repack(self):
        n = self.net
        if n.residual_conv0.conv.out_channels != 256 or n.residual_conv0.conv.in_channels != 64:
            raise NotImplementedError("Denoise: the down-sampling conv is 64 -> 256 (utils/models.py:680)")
        q4 = [slice(64 * q, 64 * q + 64) for q in range(4)]
        with torch.cuda.device(self.device):
            self.conv0 = self._pack(n.conv0[0])
            self.res0 = [[self._pack(b.m[0]), self._pack(b.m[1])] for b in n.residual_0]
            self.down = [self._pack_down(n.residual_conv0, q) for q in range(4)]
            self.res1 = [[[self._pack(b.m[j], sl) for sl in q4] for j in range(2)] for b in n.residual_1]
            self.res2 = [[self._pack(b.m[0]), self._pack(b.m[1])] for b in n.residual_2]
            self.conv1 = self._pack(n.conv1)
            self.conv2 = self._pack(n.conv2[0])
        self.versions = self._param_versions()
        self.graphs = {}

    def _blocks64(self, blocks, packs, r, keep, N, H, W):
        """x + Conv(LReLU(Conv(x))) chains on 8-plane buffers; `keep` is never written.  Returns the buffer holding the result."""
        t = self._buf("t", N, 8, H, W)
        pool = [self._buf(f"r{i}", N, 8, H, W) for i in range(2)]
        for b, pk in zip(blocks, packs):
            out = next(x for x in pool if x is not r and x is not keep)
            s0, f0 = _slope(b.m[0].act)
            ops.conv2d_fwd(pk[0], r, 8, 0, t, 8, 0, N, H, W, slope=s0, flags=f0)
            ops.conv2d_fwd(pk[1], t, 8, 0, out, 8, 0, N, H, W, residual=r, res_planes=8, res_off=0)
            r = out
        return r

    def _run_trunk(self, xin, N, H, W):
        n = self.net
        if H % 2 or W % 2:
            raise NotImplementedError(f"Denoise: H and W must be even on the sm_100a path (got {H} x {W}): the stride-2 conv runs "
                                      "over the space-to-depth input")
        h, w = H // 2, W // 2
        X0 = self._buf("x0", N, 8, H, W)                    # conv0's output: the long skip (utils/models.py:702)
        s0, f0 = _slope(n.conv0[0].act)
        ops.conv2d_fwd(self.conv0, xin, 1, 0, X0, 8, 0, N, H, W, slope=s0, flags=f0)
        r = self._blocks64(n.residual_0, self.res0, X0, X0, N, H, W)
        U = self._buf("s2d", N, 32, h, w)
        ops.pixunshuffle_mask(r, r, U, 1.0)                 # space-to-depth (slope 1: no mask)
        D = [self._buf(f"d{i}", N, 32, h, w) for i in range(3)]
        sd, fd = _slope(n.residual_conv0.act)
        for q in range(4):
            ops.conv2d_fwd(self.down[q], U, 32, 0, D[0], 32, 8 * q, N, h, w, slope=sd, flags=fd)
        d = D[0]
        for b, pk in zip(n.residual_1, self.res1):
            t = D[2]
            out = D[1] if d is D[0] else D[0]
            s1, f1 = _slope(b.m[0].act)
            for q in range(4):
                ops.conv2d_fwd(pk[0][q], d, 32, 0, t, 32, 8 * q, N, h, w, slope=s1, flags=f1)
            for q in range(4):
                ops.conv2d_fwd(pk[1][q], t, 32, 0, out, 32, 8 * q, N, h, w, residual=d, res_planes=32, res_off=8 * q)
            d = out
        up = self._buf("up", N, 8, H, W)
        ops.pixshuffle_lrelu(d, up, float(n.residual_conv1[1].negative_slope))
        r = self._blocks64(n.residual_2, self.res2, up, X0, N, H, W)
        y = self._buf("y", N, 8, H, W)
        ops.conv2d_fwd(self.conv1, r, 8, 0, y, 8, 0, N, H, W, residual=X0, res_planes=8, res_off=0)
        return y, H, W

    def _run_tail(self, cur, N, h, w, u8_out):
        tanh = ops.F_TANH if isinstance(self.net.conv2[0].act, nn.Tanh) else 0
        out = torch.empty((N, 3, h, w), dtype=torch.uint8 if u8_out else torch.float32, device=self.device)
        ops.conv2d_fwd(self.conv2, cur, 8, 0, out, 0, 0, N, h, w, flags=tanh | (ops.F_U8OUT if u8_out else ops.F_F32OUT))
        return out

    def launches(self) -> int:
        return 1 + 2 * len(self.res0) + 1 + 4 + 8 * len(self.res1) + 1 + 2 * len(self.res2) + 2


class Denoise(_SRNet):
    """utils/models.py:672-706 (current revision: 9x9 head, ResidualBlock1 stacks around a stride-2 / PixelShuffle pair).
    Inference forward on the sm_100a kernels; its training loop (train_gan.py) is outside the hot path."""
    plan_class = _DenoisePlan
    trainable = False

    def __init__(self, residual_blocks):
        super().__init__()
        self.conv0 = nn.Sequential(ConvWithoutBN(3, 64, 9, 1, act=nn.LeakyReLU(0.2)))
        self.residual_0 = nn.Sequential(*[ResidualBlock1(64, 64, 64, 3, act=nn.LeakyReLU(0.2)) for _ in range(residual_blocks // 2)])
        self.residual_conv0 = ConvWithoutBN(64, 256, 3, 2, act=nn.LeakyReLU(0.2))
        self.residual_1 = nn.Sequential(*[ResidualBlock1(256, 256, 256, 3, nn.LeakyReLU(0.2)) for _ in range(2)])
        self.residual_conv1 = nn.Sequential(*[nn.PixelShuffle(2), nn.LeakyReLU(0.2)])
        self.residual_2 = nn.Sequential(*[ResidualBlock1(64, 64, 64, 3, act=nn.LeakyReLU(0.2)) for _ in range(residual_blocks // 2)])
        self.conv1 = Conv(64, 64, 3, 1, None, act=False)
        self.conv2 = nn.Sequential(ConvWithoutBN(64, 3, 9, 1, None, act=nn.Tanh()))
        for x in self.modules():                                # utils/models.py:691-693
            if hasattr(x, "inplace"):
                x.inplace = True


class SRGAN(nn.Module):
    """utils/models.py:653-669."""

    def __init__(self, deep, add_rate, enchant=False, scaleRate=2):
        super().__init__()
        self.res_net = EResNet(deep, add_rate, scaleRate) if enchant else ResNet(deep, add_rate, scaleRate=scaleRate)

    def init_weight(self, pretrained):
        try:
            ckpt = torch.load(pretrained, "cpu", weights_only=False)
            self.res_net.load_state_dict(ckpt["ema"].float().state_dict())
            print("loaded pre-trained of Resnet")
        except Exception:  # noqa: BLE001  (the reference swallows every failure here, utils/models.py:660-665)
            print("Could not load Res checkpoint.")

    def forward(self, inputs):
        return self.res_net(inputs)


class Normalize(nn.Module):
    """utils/datasets.py:50-71 (buffers mean/std/max_pixel_value; dim=4 layout [1,3,1,1]) — parameter container;
    on the sm_100a path it runs fused into the uint8 -> C8 pack kernel."""

    def __init__(self, mean=(0.485, 0.456, 0.406), std=(0.229, 0.224, 0.225), max_pixel_value=255.0, dim=3):
        super().__init__()
        shape = (1, 3, 1, 1) if dim == 4 else (3, 1, 1)
        self.register_buffer("mean", torch.tensor(list(mean), dtype=torch.float32).reshape(shape))
        self.register_buffer("std", torch.tensor(list(std), dtype=torch.float32).reshape(shape))
        self.register_buffer("max_pixel_value", torch.tensor(float(max_pixel_value)))


class TanhToArrayImage(nn.Module):
    """utils/models.py:443-451 — fused into the tail conv's epilogue (round((t+1)/2*255) -> uint8)."""

    def __init__(self, max_pixel_value=255.0):
        super().__init__()
        self.register_buffer("max_pixel_value", torch.tensor(float(max_pixel_value)))


def fuse_conv_and_bn(conv: nn.Conv2d, bn: nn.BatchNorm2d) -> nn.Conv2d:
    """utils/models.py:366-406: W' = diag(gamma/sqrt(var+eps)) W ; b' = gamma (b - mean)/sqrt(var+eps) + beta."""
    fused = nn.Conv2d(conv.in_channels, conv.out_channels, conv.kernel_size, conv.stride, conv.padding,
                      dilation=conv.dilation, groups=conv.groups, bias=True).requires_grad_(False).to(conv.weight.device)
    scale = bn.weight / torch.sqrt(bn.running_var + bn.eps)
    with torch.no_grad():
        fused.weight.copy_(conv.weight * scale.view(-1, 1, 1, 1))
        b = conv.bias if conv.bias is not None else torch.zeros(conv.out_channels, device=conv.weight.device)
        fused.bias.copy_((b - bn.running_mean) * scale + bn.bias)
    return fused


class Model(nn.Module):
    """utils/models.py:723-761: wrapper adding the uint8 pre/post-processing and BN fusing."""
    mean = None
    std = None

    def __init__(self, model):
        super().__init__()
        self.net = model

    def init_normalize(self, mean, std):
        self.net = nn.Sequential(Normalize(mean, std, dim=4), self.net, TanhToArrayImage())

    def _inner(self):
        return self.net[1] if isinstance(self.net, nn.Sequential) else self.net

    def forward(self, inputs: torch.Tensor):
        inner = self._inner()
        core = inner.res_net if isinstance(inner, SRGAN) else inner
        if not isinstance(self.net, nn.Sequential):
            return core(inputs)
        if not inputs.is_cuda:
            raise RuntimeError("the B200 path has no CPU fallback: move the model and its inputs to a CUDA device")
        norm = self.net[0]
        if inputs.dtype != torch.uint8:
            raise NotImplementedError("Model with init_normalize takes uint8 images on the sm_100a path "
                                      "(Normalize's float branch skips the /255, utils/datasets.py:66-70)")
        import ctypes as C
        from ._lib import lib, check
        plan = core._plan(inputs.device)
        N, _, H, W = inputs.shape
        mean = (C.c_float * 3)(*norm.mean.flatten().tolist())
        std = (C.c_float * 3)(*norm.std.flatten().tolist())
        src = inputs.contiguous()

        def fill(plane):       # Normalize (utils/datasets.py:65-71) fused into the uint8 -> C8 pack, written into the plan's input plane
            check(lib().srb_u8nchw_to_c8(plan.ctx.handle, src.data_ptr(), plane.data_ptr(), mean, std,
                                         float(norm.max_pixel_value), plan.dtype, N, H, W,
                                         C.c_void_p(torch.cuda.current_stream(inputs.device).cuda_stream)), "srb_u8nchw_to_c8")
        return plan.forward(None, u8_out=True, fill_input=fill, shape=(N, 3, H, W))

    def fuse(self):
        for m in self.modules():
            if isinstance(m, Conv) and isinstance(m.bn, nn.BatchNorm2d):
                m.conv = fuse_conv_and_bn(m.conv, m.bn)
                m.fuseforward()
        return self

    def defuse(self):
        for m in self.modules():
            if isinstance(m, Conv):
                m.defuseforward()
        return self
