"""Thin Python operators over the C ABI (include/srb200.h).  torch supplies device memory and the
current stream only; all arithmetic runs in libsrb200.so's sm_100a kernels.  No fallbacks."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import (Context, check, lib, SRB_DT_F16, SRB_DT_BF16, SRB_DT_F32, F_LRELU, F_RESIDUAL, F_TANH, F_U8OUT,
                   F_PIXSHUF2, F_F32OUT)

_TORCH_DT = {SRB_DT_F16: torch.float16, SRB_DT_BF16: torch.bfloat16}


def dtype_code(dtype) -> int:
    if dtype in (SRB_DT_F16, "fp16", "float16", torch.float16):
        return SRB_DT_F16
    if dtype in (SRB_DT_BF16, "bf16", "bfloat16", torch.bfloat16):
        return SRB_DT_BF16
    raise ValueError(f"unsupported operand dtype {dtype!r} (fp16 or bf16)")


def _stream_ptr(device) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _require_cuda(t: torch.Tensor, name: str):
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor: the B200 path has no CPU fallback")
    if not t.is_contiguous():
        raise RuntimeError(f"{name} must be contiguous")


class PackedConv:
    """Packed filter of one conv (BN folded, UMMA B-operand layout) living on one device."""

    def __init__(self, ctx: Context, weight, bias=None, bn=None, bn_eps: float = 1e-5, dtype=SRB_DT_F16):
        w = np.ascontiguousarray(weight.detach().cpu().numpy() if isinstance(weight, torch.Tensor) else weight,
                                 dtype=np.float32)
        cout, cin, kh, kw = w.shape
        if kh != kw:
            raise ValueError("square filters only")
        b = None if bias is None else np.ascontiguousarray(
            bias.detach().cpu().numpy() if isinstance(bias, torch.Tensor) else bias, dtype=np.float32)
        bn4 = None
        if bn is not None:  # (gamma, beta, running_mean, running_var)
            bn4 = np.ascontiguousarray(np.stack([np.asarray(t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else t,
                                                            dtype=np.float32) for t in bn]), dtype=np.float32)
        self.ctx, self.cout, self.cin, self.k = ctx, cout, cin, kh
        self.dtype = dtype_code(dtype)
        h = C.c_void_p()
        check(lib().srb_pack_conv(ctx.handle, _lib.f32_ptr(w), None if b is None else _lib.f32_ptr(b),
                                  None if bn4 is None else _lib.f32_ptr(bn4), float(bn_eps), cout, cin, kh, self.dtype,
                                  C.byref(h)), "srb_pack_conv")
        self.handle = h

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                lib().srb_free_weights(self.handle)
                self.handle = None
        except Exception:
            pass


def conv2d_fwd(pc: PackedConv, x: torch.Tensor, x_planes: int, x_off: int, y: torch.Tensor, y_planes: int, y_off: int,
               N: int, H: int, W: int, residual: torch.Tensor | None = None, res_planes: int = 0, res_off: int = 0,
               res_scale: float = 1.0, out_scale: float = 1.0, slope: float = 0.0, flags: int = 0,
               res2: torch.Tensor | None = None, res2_planes: int = 0, res2_off: int = 0, res2_scale: float = 1.0):
    _require_cuda(x, "x")
    _require_cuda(y, "y")
    if residual is not None:
        _require_cuda(residual, "residual")
        flags |= F_RESIDUAL
    if res2 is not None:
        _require_cuda(res2, "res2")
    check(lib().srb_conv2d_fwd_res2(pc.ctx.handle, pc.handle, x.data_ptr(), x_planes, x_off, y.data_ptr(), y_planes, y_off,
                                    None if residual is None else residual.data_ptr(), res_planes, res_off, res_scale,
                                    None if res2 is None else res2.data_ptr(), res2_planes, res2_off, res2_scale,
                                    out_scale, slope, flags, N, H, W, _stream_ptr(x.device)), "srb_conv2d_fwd")


def c8_empty(N: int, planes: int, H: int, W: int, dtype, device) -> torch.Tensor:
    """[N][planes][H][W][8] activation buffer."""
    return torch.empty((N, planes, H, W, 8), dtype=_TORCH_DT[dtype_code(dtype)], device=device)


def nchw_to_c8(x: torch.Tensor, out: torch.Tensor | None = None, planes: int | None = None, off: int = 0,
               dtype=SRB_DT_F16) -> torch.Tensor:
    """fp32 NCHW -> C8 planes (zero padded to a multiple of 8 channels) with the library's kernel."""
    _require_cuda(x, "x")
    x = x.float().contiguous()
    N, Cc, H, W = x.shape
    p = (Cc + 7) // 8
    planes = planes or p
    if out is None:
        out = c8_empty(N, planes, H, W, dtype, x.device)
    ctx = Context.get(x.device.index or 0)
    check(lib().srb_f32nchw_to_c8(ctx.handle, x.data_ptr(), out.data_ptr(), Cc, planes, off, dtype_code(dtype), N, H, W,
                                  _stream_ptr(x.device)), "srb_f32nchw_to_c8")
    return out


def c8_to_nchw(x: torch.Tensor, channels: int, off: int = 0) -> torch.Tensor:
    _require_cuda(x, "x")
    N, planes, H, W, _ = x.shape
    out = torch.empty((N, channels, H, W), dtype=torch.float32, device=x.device)
    ctx = Context.get(x.device.index or 0)
    check(lib().srb_c8_to_f32nchw(ctx.handle, x.data_ptr(), out.data_ptr(), channels, planes, off,
                                  dtype_code(x.dtype), N, H, W, _stream_ptr(x.device)), "srb_c8_to_f32nchw")
    return out


def _pred_code(t: torch.Tensor) -> int:
    return {torch.float16: SRB_DT_F16, torch.bfloat16: SRB_DT_BF16, torch.float32: SRB_DT_F32}[t.dtype]


def l1_loss_fwd(pred: torch.Tensor, tgt: torch.Tensor, w: float = 1.0, squared: bool = False) -> torch.Tensor:
    """w * mean|pred - tgt| (squared: w * mean (pred - tgt)^2, nn.MSELoss)."""
    _require_cuda(pred, "pred")
    _require_cuda(tgt, "tgt")
    if tgt.dtype != torch.float32 or pred.shape != tgt.shape:
        raise RuntimeError("l1_loss_fwd: target must be fp32 with pred's shape")
    ctx = Context.get(pred.device.index or 0)
    out = torch.empty((), dtype=torch.float32, device=pred.device)
    partials = torch.empty(lib().srb_l1_partials_len(), dtype=torch.float32, device=pred.device)
    fn = lib().srb_mse_loss_fwd if squared else lib().srb_l1_loss_fwd
    check(fn(ctx.handle, pred.data_ptr(), _pred_code(pred), tgt.data_ptr(), pred.numel(), float(w),
             out.data_ptr(), partials.data_ptr(), _stream_ptr(pred.device)), "srb_mse_loss_fwd" if squared else "srb_l1_loss_fwd")
    return out


def l1_loss_bwd(pred: torch.Tensor, tgt: torch.Tensor, w: float = 1.0, grad_scale: float = 1.0) -> torch.Tensor:
    _require_cuda(pred, "pred")
    _require_cuda(tgt, "tgt")
    ctx = Context.get(pred.device.index or 0)
    g = torch.empty_like(pred)
    check(lib().srb_l1_loss_bwd(ctx.handle, pred.data_ptr(), _pred_code(pred), tgt.data_ptr(), pred.numel(), float(w),
                                float(grad_scale), g.data_ptr(), _stream_ptr(pred.device)), "srb_l1_loss_bwd")
    return g


def conv2d_wgrad(x: torch.Tensor, x_planes: int, x_off: int, cin: int, dy: torch.Tensor, dy_planes: int, dy_off: int,
                 cout: int, dw: torch.Tensor, N: int, H: int, W: int, scale: float = 1.0, k: int = 3):
    """dw (fp32 OIHW, device) += scale * wgrad(x slice, dy slice)."""
    _require_cuda(x, "x"); _require_cuda(dy, "dy"); _require_cuda(dw, "dw")
    if dw.dtype != torch.float32 or tuple(dw.shape) != (cout, cin, k, k):
        raise RuntimeError("conv2d_wgrad: dw must be fp32 [cout, cin, k, k]")
    ctx = Context.get(x.device.index or 0)
    check(lib().srb_conv2d_wgrad(ctx.handle, x.data_ptr(), x_planes, x_off, cin, dy.data_ptr(), dy_planes, dy_off, cout,
                                 dw.data_ptr(), float(scale), k, dtype_code(x.dtype), N, H, W, _stream_ptr(x.device)),
          "srb_conv2d_wgrad")


def bias_grad(dy: torch.Tensor, dy_planes: int, dy_off: int, cout: int, db: torch.Tensor, N: int, H: int, W: int,
              scale: float = 1.0):
    _require_cuda(dy, "dy"); _require_cuda(db, "db")
    ctx = Context.get(dy.device.index or 0)
    check(lib().srb_bias_grad(ctx.handle, dy.data_ptr(), dy_planes, dy_off, cout, db.data_ptr(), float(scale),
                              dtype_code(dy.dtype), N, H, W, _stream_ptr(dy.device)), "srb_bias_grad")


# ---------------------------------------------------------------------------------------------
# training-step operators (device-resident fp32 master weights, backward glue)
# ---------------------------------------------------------------------------------------------
class DeviceConv:
    """Packed operand of a (cout, cin, k) conv that is (re)filled on the device from fp32 master weights."""

    def __init__(self, ctx: Context, cout: int, cin: int, k: int, dtype, ws: int = 0):
        """ws = 1: weight-streaming layout also for cin = 64 (the conv can then be a layer of a ConvChain);
        ws = 2: ... with the dy taps folded into N (MODE_DYF: one N = 96 MMA per three taps; 32 outputs)."""
        self.ctx, self.cout, self.cin, self.k = ctx, cout, cin, k
        self.dtype = dtype_code(dtype)
        h = C.c_void_p()
        check(lib().srb_alloc_conv_ex(ctx.handle, cout, cin, k, self.dtype, int(ws), C.byref(h)), "srb_alloc_conv_ex")
        self.handle = h

    def fill(self, weight: torch.Tensor, bias: torch.Tensor | None = None, o_off: int = 0, c_off: int = 0,
             transpose: bool = False):
        """weight: fp32 [O][I][k][k] on the device.  transpose=True packs the data-gradient operand."""
        _require_cuda(weight, "weight")
        if weight.dtype != torch.float32:
            raise RuntimeError("master weights must be fp32")
        check(lib().srb_pack_conv_device(self.ctx.handle, self.handle, weight.data_ptr(), weight.shape[0], weight.shape[1],
                                         o_off, c_off, 1 if transpose else 0,
                                         None if (bias is None or transpose) else bias.data_ptr(),
                                         _stream_ptr(weight.device)), "srb_pack_conv_device")

    __del__ = PackedConv.__del__


def conv2d_ex(pc, x, x_planes, x_off, y, y_planes, y_off, N, H, W, residual=None, res_planes=0, res_off=0, res_nplanes=0,
              res_scale=1.0, res2=None, res2_planes=0, res2_off=0, res2_scale=1.0, out_scale=1.0, slope=0.0, flags=0,
              mask=None, mask_planes=0, mask_off=0, mask_p0=0, mask_p1=0, mask_slope=0.0):
    """General conv launch (srb_conv2d_ex): forward epilogues plus the backward extras (partial residual, mask)."""
    a = _lib.ConvArgs()
    a.x, a.x_c8_total, a.x_c8_off = x.data_ptr(), x_planes, x_off
    a.y, a.y_c8_total, a.y_c8_off = y.data_ptr(), y_planes, y_off
    if residual is not None:
        flags |= F_RESIDUAL
        a.residual, a.res_c8_total, a.res_c8_off, a.res_nplanes = residual.data_ptr(), res_planes, res_off, res_nplanes
    a.res_scale = res_scale
    if res2 is not None:
        a.res2, a.res2_c8_total, a.res2_c8_off, a.res2_scale = res2.data_ptr(), res2_planes, res2_off, res2_scale
    a.out_scale, a.lrelu_slope = out_scale, slope
    if mask is not None:
        flags |= _lib.F_MASK
        a.mask, a.mask_c8_total, a.mask_c8_off = mask.data_ptr(), mask_planes, mask_off
        a.mask_p0, a.mask_p1, a.mask_slope = mask_p0, mask_p1, mask_slope
    a.flags, a.N, a.H, a.W = flags, N, H, W
    check(lib().srb_conv2d_ex(pc.ctx.handle, pc.handle, C.byref(a), _stream_ptr(x.device)), "srb_conv2d_ex")


def conv_args(x, x_planes, x_off, y, y_planes, y_off, N, H, W, slope=0.0, flags=0, out_scale=1.0,
              mask=None, mask_planes=0, mask_off=0, mask_p0=0, mask_p1=0, mask_slope=0.0,
              residual=None, res_planes=0, res_off=0, res_scale=1.0, res2=None, res2_planes=0, res2_off=0, res2_scale=1.0) -> "_lib.ConvArgs":
    """srb_conv_args of a layer of a ConvChain (residual terms: the block's last conv)."""
    a = _lib.ConvArgs()
    a.x, a.x_c8_total, a.x_c8_off = x.data_ptr(), x_planes, x_off
    a.y, a.y_c8_total, a.y_c8_off = y.data_ptr(), y_planes, y_off
    a.res_scale, a.out_scale, a.lrelu_slope = res_scale, out_scale, slope
    if residual is not None:
        flags |= _lib.F_RESIDUAL
        a.residual, a.res_c8_total, a.res_c8_off = residual.data_ptr(), res_planes, res_off
    if res2 is not None:
        a.res2, a.res2_c8_total, a.res2_c8_off, a.res2_scale = res2.data_ptr(), res2_planes, res2_off, res2_scale
    if mask is not None:
        flags |= _lib.F_MASK
        a.mask, a.mask_c8_total, a.mask_c8_off = mask.data_ptr(), mask_planes, mask_off
        a.mask_p0, a.mask_p1, a.mask_slope = mask_p0, mask_p1, mask_slope
    a.flags, a.N, a.H, a.W = flags, N, H, W
    return a


class BatchNormPass:
    """The four train-mode BatchNorm passes of one `Conv` block (srb_bn_*; utils/models.py:75-111) over fixed buffers: the
    argument block is built once, `stats / apply / bwd_reduce / bwd_apply` launch on the current stream."""

    def __init__(self, ctx: Context, dtype, N, H, W, planes, y, y_planes, y_off, gamma, beta, sums, bsums, bn=None, slope=0.0,
                 act=0, out=None, out_planes=0, out_off=0, out_scale=1.0, res=None, res_planes=0, res_off=0, res_scale=1.0,
                 res2=None, res2_planes=0, res2_off=0, res2_scale=1.0):
        self.ctx, self.dt = ctx, dtype_code(dtype)
        a = self.a = _lib.BNArgs()
        a.y, a.y_c8_total, a.y_c8_off = y.data_ptr(), y_planes, y_off
        a.planes, a.gamma, a.beta = planes, gamma.data_ptr(), beta.data_ptr()
        a.sums, a.bsums = sums.data_ptr(), bsums.data_ptr()
        if bn is not None and bn.running_mean is not None:
            a.running_mean, a.running_var = bn.running_mean.data_ptr(), bn.running_var.data_ptr()
            a.num_batches_tracked = bn.num_batches_tracked.data_ptr()
            a.momentum = 0.1 if bn.momentum is None else float(bn.momentum)
        a.eps = 1e-5 if bn is None else float(bn.eps)
        a.slope, a.act, a.out_scale = float(slope), int(act), float(out_scale)
        if out is not None:
            a.out, a.out_c8_total, a.out_c8_off = out.data_ptr(), out_planes, out_off
        if res is not None:
            a.res, a.res_c8_total, a.res_c8_off, a.res_scale = res.data_ptr(), res_planes, res_off, float(res_scale)
        if res2 is not None:
            a.res2, a.res2_c8_total, a.res2_c8_off, a.res2_scale = res2.data_ptr(), res2_planes, res2_off, float(res2_scale)
        a.up_scale, a.grad_scale = 1.0, 1.0
        a.N, a.H, a.W = N, H, W
        self._keep = (y, gamma, beta, sums, bsums, bn, out, res, res2)
        self._fwd = C.create_string_buffer(bytes(a), C.sizeof(a))       # the forward's argument block (backward re-targets `out`)

    def _run(self, fn, name, args, device):
        check(fn(self.ctx.handle, C.byref(args), self.dt, _stream_ptr(device)), name)

    def stats(self, device):
        self._run(lib().srb_bn_stats, "srb_bn_stats", self.a, device)

    def apply(self, device):
        self._run(lib().srb_bn_apply, "srb_bn_apply", self.a, device)

    def set_backward(self, da, da_planes, da_off, out, out_planes, out_off, dgamma, dbeta, up_scale=1.0, grad_scale=1.0):
        b = self.b = _lib.BNArgs.from_buffer_copy(self._fwd)
        b.da, b.da_c8_total, b.da_c8_off, b.up_scale = da.data_ptr(), da_planes, da_off, float(up_scale)
        b.out, b.out_c8_total, b.out_c8_off = out.data_ptr(), out_planes, out_off
        b.dgamma, b.dbeta, b.grad_scale = dgamma.data_ptr(), dbeta.data_ptr(), float(grad_scale)
        self._keep_b = (da, out, dgamma, dbeta)

    def bwd_reduce(self, device):
        self._run(lib().srb_bn_bwd_reduce, "srb_bn_bwd_reduce", self.b, device)

    def bwd_apply(self, device):
        self._run(lib().srb_bn_bwd_apply, "srb_bn_bwd_apply", self.b, device)


class ConvChain:
    """Up to four weight-streaming 3x3 convs (32 outputs each) of one dense block as ONE launch (srb_conv_chain_*):
    layers = [(packed conv, ConvArgs)]; the tensors named by the arguments must outlive the chain."""

    def __init__(self, ctx: Context, layers: list, keep=()):
        self.ctx, self._keep = ctx, (layers, keep)
        n = len(layers)
        wp = (C.c_void_p * n)(*[pc.handle for pc, _ in layers])
        arr = (_lib.ConvArgs * n)()
        for i, (_, a) in enumerate(layers):
            C.memmove(C.byref(arr[i]), C.byref(a), C.sizeof(_lib.ConvArgs))
        h = C.c_void_p()
        check(lib().srb_conv_chain_create(ctx.handle, n, wp, C.cast(arr, C.c_void_p), C.byref(h)), "srb_conv_chain_create")
        self.handle = h

    def run(self, device):
        check(lib().srb_conv_chain_run(self.ctx.handle, self.handle, _stream_ptr(device)), "srb_conv_chain_run")

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                lib().srb_conv_chain_destroy(self.handle)
                self.handle = None
        except Exception:
            pass


def conv9_wgrad(x, x_planes, x_off, cin, dy, dy_planes, dy_off, cout, dw, N, H, W, scale=1.0):
    _require_cuda(x, "x"); _require_cuda(dy, "dy"); _require_cuda(dw, "dw")
    if dw.dtype != torch.float32 or dw.numel() != cout * cin * 81:
        raise RuntimeError("conv9_wgrad: dw must be fp32 [cout, cin, 9, 9]")
    ctx = Context.get(x.device.index or 0)
    check(lib().srb_conv9_wgrad(ctx.handle, x.data_ptr(), x_planes, x_off, cin, dy.data_ptr(), dy_planes, dy_off, cout,
                                dw.data_ptr(), float(scale), dtype_code(x.dtype), N, H, W, _stream_ptr(x.device)),
          "srb_conv9_wgrad")


def tanh_bwd_to_c8(gout: torch.Tensor, out: torch.Tensor, dtype, scale: float = 1.0, g8: torch.Tensor | None = None) -> torch.Tensor:
    _require_cuda(gout, "gout"); _require_cuda(out, "out")
    N, Cc, H, W = out.shape
    if Cc != 3 or gout.shape != out.shape or gout.dtype != torch.float32 or out.dtype != torch.float32:
        raise RuntimeError("tanh_bwd_to_c8: fp32 [N,3,H,W] tensors expected")
    if g8 is None:
        g8 = c8_empty(N, 1, H, W, dtype, out.device)
    ctx = Context.get(out.device.index or 0)
    check(lib().srb_tanh_bwd_to_c8(ctx.handle, gout.data_ptr(), out.data_ptr(), g8.data_ptr(), float(scale), dtype_code(dtype),
                                   N, H, W, _stream_ptr(out.device)), "srb_tanh_bwd_to_c8")
    return g8


def pixunshuffle_mask(g_u: torch.Tensor, u: torch.Tensor, g_s: torch.Tensor, slope: float):
    """g_u, u: [N][p][2H][2W][8]; g_s: [N][4p][H][W][8]."""
    N, p, H2, W2, _ = g_u.shape
    ctx = Context.get(g_u.device.index or 0)
    check(lib().srb_pixunshuffle_mask(ctx.handle, g_u.data_ptr(), u.data_ptr(), g_s.data_ptr(), p, float(slope),
                                      dtype_code(g_u.dtype), N, H2 // 2, W2 // 2, _stream_ptr(g_u.device)), "srb_pixunshuffle_mask")


def pixshuffle_lrelu(s: torch.Tensor, u: torch.Tensor, slope: float):
    """s: [N][4p][H][W][8] -> u: [N][p][2H][2W][8] = LeakyReLU(PixelShuffle(2)(s))."""
    N, p, H2, W2, _ = u.shape
    ctx = Context.get(u.device.index or 0)
    check(lib().srb_pixshuffle_lrelu(ctx.handle, s.data_ptr(), u.data_ptr(), p, float(slope), dtype_code(u.dtype), N, H2 // 2, W2 // 2,
                                     _stream_ptr(u.device)), "srb_pixshuffle_lrelu")


def c8_axpby_mask(a, a_planes, a_off, alpha, b, b_planes, b_off, beta, act, act_planes, act_off, slope, out, out_planes, out_off,
                  planes, N, H, W):
    ctx = Context.get(a.device.index or 0)
    check(lib().srb_c8_axpby_mask(ctx.handle, a.data_ptr(), a_planes, a_off, float(alpha),
                                  None if b is None else b.data_ptr(), b_planes, b_off, float(beta),
                                  None if act is None else act.data_ptr(), act_planes, act_off, float(slope),
                                  out.data_ptr(), out_planes, out_off, planes, dtype_code(a.dtype), N, H, W,
                                  _stream_ptr(a.device)), "srb_c8_axpby_mask")


def l1_loss_bwd_dev(pred: torch.Tensor, tgt: torch.Tensor, grad_out: torch.Tensor, w: float = 1.0, squared: bool = False) -> torch.Tensor:
    """dpred = grad_out (device scalar) * w * sign(pred - tgt) / numel  (squared: ... * 2 (pred - tgt) / numel)."""
    _require_cuda(pred, "pred"); _require_cuda(tgt, "tgt"); _require_cuda(grad_out, "grad_out")
    ctx = Context.get(pred.device.index or 0)
    g = torch.empty_like(pred)
    go = grad_out.detach().reshape(()).float().contiguous()
    fn = lib().srb_mse_loss_bwd_dev if squared else lib().srb_l1_loss_bwd_dev
    check(fn(ctx.handle, pred.data_ptr(), _pred_code(pred), tgt.data_ptr(), pred.numel(), float(w), 1.0,
             go.data_ptr(), g.data_ptr(), _stream_ptr(pred.device)), "srb_mse_loss_bwd_dev" if squared else "srb_l1_loss_bwd_dev")
    return g


class GradBatch:
    """All 3x3 weight gradients + bias gradients of one backward pass: one launch each (srb_grad_batch_*).
    jobs: dicts with x, x_planes, x_off, cin, dy, dy_planes, dy_off, cout, dw, db, scale, N, H, W (x/dw or db may be None)."""

    def __init__(self, ctx: Context, jobs: list[dict], dtype):
        self.ctx = ctx
        self._keep = jobs                      # the tensors named by the jobs must outlive the batch
        arr = (_lib.GradJob * len(jobs))()
        for a, j in zip(arr, jobs):
            a.x = None if j.get("x") is None else j["x"].data_ptr()
            a.x_c8_total, a.x_c8_off, a.cin = j.get("x_planes", 0), j.get("x_off", 0), j.get("cin", 0)
            a.dy, a.dy_c8_total, a.dy_c8_off, a.cout = j["dy"].data_ptr(), j["dy_planes"], j["dy_off"], j["cout"]
            a.dw = None if j.get("dw") is None else j["dw"].data_ptr()
            a.db = None if j.get("db") is None else j["db"].data_ptr()
            a.scale, a.N, a.H, a.W = float(j.get("scale", 1.0)), j["N"], j["H"], j["W"]
        h = C.c_void_p()
        check(lib().srb_grad_batch_create(ctx.handle, arr, len(jobs), dtype_code(dtype), C.byref(h)), "srb_grad_batch_create")
        self.handle = h
        self.units = lib().srb_grad_batch_units(h)

    def run(self, device, splitk: int = 1):
        check(lib().srb_grad_batch_run(self.ctx.handle, self.handle, splitk, _stream_ptr(device)), "srb_grad_batch_run")

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                lib().srb_grad_batch_destroy(self.handle)
                self.handle = None
        except Exception:
            pass


class PackBatch:
    """All packed operands of a net re-derived from the fp32 master weights in one launch (srb_pack_batch_*).
    jobs: (DeviceConv, weight, bias_or_None, o_off, c_off, transpose[, c_lo, c_n, scale]) — the optional tail fills only the
    operand's input channels [c_lo, c_lo + c_n) (gathered operands: one job per source filter) with scaled weights."""

    def __init__(self, ctx: Context, jobs: list[tuple]):
        self.ctx = ctx
        self._keep = jobs
        arr = (_lib.PackJob * len(jobs))()
        for a, job in zip(arr, jobs):
            pc, w, b, o_off, c_off, tr = job[:6]
            a.c_lo, a.c_n, a.scale = (job[6], job[7], float(job[8])) if len(job) > 6 else (0, 0, 1.0)
            if w.dtype != torch.float32 or not w.is_cuda or not w.is_contiguous():
                raise RuntimeError("master weights must be contiguous fp32 CUDA tensors")
            a.w, a.w_src, a.src_o, a.src_i = pc.handle, w.data_ptr(), w.shape[0], w.shape[1]
            a.o_off, a.c_off, a.transpose = o_off, c_off, 1 if tr else 0
            a.bias_src = None if (b is None or tr) else b.data_ptr()
        h = C.c_void_p()
        check(lib().srb_pack_batch_create(ctx.handle, arr, len(jobs), C.byref(h)), "srb_pack_batch_create")
        self.handle = h

    def run(self, device):
        check(lib().srb_pack_batch_run(self.ctx.handle, self.handle, _stream_ptr(device)), "srb_pack_batch_run")

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                lib().srb_pack_batch_destroy(self.handle)
                self.handle = None
        except Exception:
            pass


class RowConv:
    """Row-fold 3x3 conv 64 -> 64 (+ bias, optional LeakyReLU, optional residual) in one launch (srb_rowconv_*):
    `Conv(64,64,3,act=False)` + long skip of the model.pt graph (utils/models.py:690, :704)."""

    def __init__(self, ctx: Context, w, b, dtype=SRB_DT_F16):
        arrs = [np.ascontiguousarray(t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else t, dtype=np.float32)
                for t in (w, b)]
        if arrs[0].shape != (64, 64, 3, 3) or arrs[1].shape != (64,):
            raise ValueError("RowConv: weight must be [64][64][3][3], bias [64]")
        self.ctx, self.dtype = ctx, dtype_code(dtype)
        h = C.c_void_p()
        check(lib().srb_rowconv_create(ctx.handle, *[_lib.f32_ptr(a) for a in arrs], self.dtype, C.byref(h)), "srb_rowconv_create")
        self.handle = h

    def forward(self, x: torch.Tensor, y: torch.Tensor, res: torch.Tensor | None = None, slope: float = 0.0, lrelu: bool = False):
        _require_cuda(x, "x"); _require_cuda(y, "y")
        N, planes, H, W, _ = x.shape
        if planes != 8 or y.shape != x.shape or (res is not None and res.shape != x.shape):
            raise RuntimeError("RowConv: x, y (and res) must be [N][8][H][W][8]")
        if res is not None:
            _require_cuda(res, "res")
        check(lib().srb_rowconv_fwd(self.ctx.handle, self.handle, x.data_ptr(), res.data_ptr() if res is not None else None,
                                    y.data_ptr(), float(slope), F_LRELU if lrelu else 0, N, H, W, _stream_ptr(x.device)),
              "srb_rowconv_fwd")
        return y

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                lib().srb_rowconv_destroy(self.handle)
                self.handle = None
        except Exception:
            pass


class Head:
    """Normalize + conv0 9x9 3 -> 64 (+ bias, LeakyReLU) in one launch straight from uint8 frames (srb_head_*): `norm_(frames)`
    (rs.py:62-63, utils/datasets.py:65-71) fused into the first conv of the model.pt graph (utils/models.py:687, :97-98)."""
    LAYOUTS = {"nchw": 0, "hwc_bgr": 1, "hwc_rgb": 2}

    def __init__(self, ctx: Context, w, b, dtype=SRB_DT_F16):
        arrs = [np.ascontiguousarray(t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else t, dtype=np.float32)
                for t in (w, b)]
        if arrs[0].shape != (64, 3, 9, 9) or arrs[1].shape != (64,):
            raise ValueError("Head: weight must be [64][3][9][9], bias [64]")
        self.ctx, self.dtype = ctx, dtype_code(dtype)
        h = C.c_void_p()
        check(lib().srb_head_create(ctx.handle, *[_lib.f32_ptr(a) for a in arrs], self.dtype, C.byref(h)), "srb_head_create")
        self.handle = h

    def forward(self, x: torch.Tensor, mean, std, slope: float, layout: str = "nchw", max_pixel_value: float = 255.0,
                out: torch.Tensor | None = None) -> torch.Tensor:
        _require_cuda(x, "x")
        if x.dtype != torch.uint8 or x.dim() != 4 or not x.is_contiguous():
            raise RuntimeError("Head: x must be a contiguous uint8 tensor [N,3,H,W] (nchw) or [N,H,W,3] (hwc_*)")
        if layout == "nchw":
            N, ch, H, W = x.shape
        else:
            N, H, W, ch = x.shape
        if ch != 3:
            raise RuntimeError("Head: three channels")
        y = out if out is not None else c8_empty(N, 8, H, W, torch.bfloat16 if self.dtype == SRB_DT_BF16 else torch.float16, x.device)
        check(lib().srb_head_fwd(self.ctx.handle, self.handle, x.data_ptr(), self.LAYOUTS[layout], (C.c_float * 3)(*mean),
                                 (C.c_float * 3)(*std), float(max_pixel_value), y.data_ptr(), float(slope), N, H, W,
                                 _stream_ptr(x.device)), "srb_head_fwd")
        return y

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                lib().srb_head_destroy(self.handle)
                self.handle = None
        except Exception:
            pass


class Tail9:
    """Row-fold 9x9 conv 64 -> 3 + bias (+ tanh) -> NCHW uint8 / fp32 in one launch (srb_tail9_*): conv2 + TanhToArrayImage
    of the model.pt graph (utils/models.py:691, :448-451)."""

    def __init__(self, ctx: Context, w, b, dtype=SRB_DT_F16):
        arrs = [np.ascontiguousarray(t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else t, dtype=np.float32)
                for t in (w, b)]
        if arrs[0].shape != (3, 64, 9, 9) or arrs[1].shape != (3,):
            raise ValueError("Tail9: weight must be [3][64][9][9], bias [3]")
        self.ctx, self.dtype = ctx, dtype_code(dtype)
        h = C.c_void_p()
        check(lib().srb_tail9_create(ctx.handle, *[_lib.f32_ptr(a) for a in arrs], self.dtype, C.byref(h)), "srb_tail9_create")
        self.handle = h

    def forward(self, x: torch.Tensor, tanh: bool = True, u8: bool = True, out: torch.Tensor = None) -> torch.Tensor:
        _require_cuda(x, "x")
        N, planes, H, W, _ = x.shape
        if planes != 8:
            raise RuntimeError("Tail9: x must be [N][8][H][W][8]")
        dt = torch.uint8 if u8 else torch.float32
        if out is None:
            out = torch.empty(N, 3, H, W, dtype=dt, device=x.device)
        elif out.dtype != dt or tuple(out.shape) != (N, 3, H, W) or not out.is_contiguous() or out.device != x.device:
            raise RuntimeError("Tail9: out must be a contiguous [N,3,H,W] tensor of the output type on x's device")
        y = out
        flags = (F_TANH if tanh else 0) | (F_U8OUT if u8 else F_F32OUT)
        check(lib().srb_tail9_fwd(self.ctx.handle, self.handle, x.data_ptr(), y.data_ptr(), flags, N, H, W, _stream_ptr(x.device)),
              "srb_tail9_fwd")
        return y

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                lib().srb_tail9_destroy(self.handle)
                self.handle = None
        except Exception:
            pass


def c1tail_supported(H: int, W: int) -> bool:
    return bool(lib().srb_c1tail_supported(H, W))


def c1tail_forward(rc: "RowConv", t: "Tail9", x: torch.Tensor, skip: torch.Tensor, tanh: bool = True, u8: bool = True,
                   out: torch.Tensor | None = None, hwc_bgr: bool = False) -> torch.Tensor:
    """conv1 + long skip + conv2 9x9 (+ tanh) -> uint8 / fp32 in ONE launch (srb_c1tail_fwd): `rc.forward(x, tmp, skip)` followed
    by `t.forward(tmp)` with tmp kept in shared memory (utils/models.py:690, :704, :691, :448-451); bit-identical to the two calls."""
    _require_cuda(x, "x"); _require_cuda(skip, "skip")
    N, planes, H, W, _ = x.shape
    if planes != 8 or skip.shape != x.shape:
        raise RuntimeError("c1tail_forward: x and skip must be [N][8][H][W][8]")
    dt = torch.uint8 if u8 else torch.float32
    oshape = (N, H, W, 3) if hwc_bgr else (N, 3, H, W)
    if out is None:
        out = torch.empty(oshape, dtype=dt, device=x.device)
    elif out.dtype != dt or out.numel() != N * 3 * H * W or not out.is_contiguous() or out.device != x.device:
        raise RuntimeError("c1tail_forward: out must be a contiguous tensor of the output type and size on x's device")
    flags = (F_TANH if tanh else 0) | (F_U8OUT if u8 else F_F32OUT) | (128 if hwc_bgr else 0)
    check(lib().srb_c1tail_fwd(rc.ctx.handle, rc.handle, t.handle, x.data_ptr(), skip.data_ptr(), out.data_ptr(), flags, N, H, W,
                               _stream_ptr(x.device)), "srb_c1tail_fwd")
    return out


class ResBlock:
    """Fused ResidualBlock1 (utils/models.py:208-209): y = x + convB(LeakyReLU(convA(x))) in one kernel (srb_resblock_*)."""

    def __init__(self, ctx: Context, wA, bA, wB, bB, dtype=SRB_DT_F16):
        arrs = [np.ascontiguousarray(t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else t, dtype=np.float32)
                for t in (wA, bA, wB, bB)]
        if arrs[0].shape != (32, 64, 3, 3) or arrs[2].shape != (64, 32, 3, 3):
            raise ValueError("ResBlock: convA must be 64->32 3x3 and convB 32->64 3x3")
        self.ctx, self.dtype = ctx, dtype_code(dtype)
        h = C.c_void_p()
        check(lib().srb_resblock_create(ctx.handle, *[_lib.f32_ptr(a) for a in arrs], self.dtype, C.byref(h)), "srb_resblock_create")
        self.handle = h

    @staticmethod
    def supported(H: int, W: int) -> bool:
        return bool(lib().srb_resblock_supported(H, W))

    def forward(self, x: torch.Tensor, y: torch.Tensor, slope: float):
        _require_cuda(x, "x"); _require_cuda(y, "y")
        N, planes, H, W, _ = x.shape
        if planes != 8 or y.shape != x.shape:
            raise RuntimeError("ResBlock: x and y must be [N][8][H][W][8]")
        check(lib().srb_resblock_fwd(self.ctx.handle, self.handle, x.data_ptr(), y.data_ptr(), float(slope), N, H, W,
                                     _stream_ptr(x.device)), "srb_resblock_fwd")
        return y

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                lib().srb_resblock_destroy(self.handle)
                self.handle = None
        except Exception:
            pass
