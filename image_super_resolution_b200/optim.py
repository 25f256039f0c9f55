"""Fused optimizer tail of the training step: drop-ins for the objects train.py:58-63 calls —
`gradscaler.unscale_(optimizer)`, `clip_grad_norm_(model.parameters(), 10)`, `gradscaler.step(optimizer)`,
`gradscaler.update()`, `ema.update(model)` — that keep the reference's call order and signatures but run the work as TWO
launches over flat fp32 arenas (libsrb200: srb_grad_stats, srb_adam_ema_step) instead of ~2 000 per-tensor launches:

  FlatAdam(params, lr, betas, eps, weight_decay)   torch.optim.Adam's signature and arithmetic (train.py:265); moves the
                                                   parameters into one flat arena (views keep their names / shapes), keeps
                                                   exp_avg / exp_avg_sq flat, reads the gradients where backward wrote them
                                                   (the flat gradient arena of training.TrainPlan — also what DDP all-reduces)
  GradScaler(...)                                  torch.amp.GradScaler whose unscale_ / step recognise a FlatAdam: the
                                                   inf/nan check and the global norm are one read of the gradient arena, the
                                                   step is skipped ON THE DEVICE when a gradient is non-finite (no .item())
  clip_grad_norm_(parameters, max_norm)            torch's signature; for a FlatAdam's parameters it only records max_norm —
                                                   the coefficient is applied where the Adam moments read the gradient
  ModelEMA.attach(optimizer)  (train.py)           the EMA blend rides on the Adam pass (same formula, utils/models.py:31-40)

Everything else (LinearLR, zero_grad, state_dict / load_state_dict) is torch.optim.Optimizer's.  Other optimizers and
parameter lists fall through to torch's own implementations.  No CPU path: FlatAdam refuses CPU parameters.
"""
from __future__ import annotations

import ctypes as C
import weakref

import torch
from torch.amp.grad_scaler import OptState

from . import _lib
from ._lib import Context, lib, check

_OWNER: "weakref.WeakValueDictionary[int, FlatAdam]" = weakref.WeakValueDictionary()    # id(parameter) -> its FlatAdam


def _stream_ptr(device) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def flatten_(tensors: list[torch.Tensor]) -> torch.Tensor:
    """Move `tensors` (same device, fp32) into one zero-padded flat arena; every tensor becomes a view of it (same shape,
    same values).  Returns the arena (length a multiple of 4 floats)."""
    dev = tensors[0].device
    n = sum(t.numel() for t in tensors)
    arena = torch.zeros((n + 3) // 4 * 4, dtype=torch.float32, device=dev)
    o = 0
    for t in tensors:
        if t.device != dev or t.dtype != torch.float32:
            raise RuntimeError("flat arenas hold fp32 tensors of one CUDA device")
        v = arena[o:o + t.numel()].view(t.shape)
        v.copy_(t.data)
        t.data = v
        o += t.numel()
    return arena


class FlatAdam(torch.optim.Optimizer):
    """torch.optim.Adam (no amsgrad / maximize) over flat arenas; see the module docstring."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        if len(self.param_groups) != 1:
            raise NotImplementedError("FlatAdam: one parameter group (train.py:265 builds Adam(model.parameters(), lr))")
        ps = self.param_groups[0]["params"]
        if not ps or any(not p.is_cuda for p in ps):
            raise RuntimeError("FlatAdam: parameters must live on a CUDA device (model.to(device) first); there is no CPU path")
        self.device = ps[0].device
        self.ctx = Context.get(self.device.index or 0)
        with torch.no_grad():
            self.p_arena = flatten_(ps)
        self.n = self.p_arena.numel()
        self.m_arena = torch.zeros_like(self.p_arena)
        self.v_arena = torch.zeros_like(self.p_arena)
        # stats: [sum (g/scale)^2, found_inf, steps taken, -]; workspace of the norm reduction (zeroed once)
        self.stats = torch.zeros(4, dtype=torch.float32, device=self.device)
        self.work = torch.zeros(lib().srb_optim_workspace_floats(), dtype=torch.float32, device=self.device)
        self.offsets, o = [], 0
        for p in ps:
            self.offsets.append(o)
            # torch.optim.Adam's per-parameter state layout (state_dict interchange); `step` is one shared device scalar
            self.state[p] = dict(step=self.stats[2], exp_avg=self.m_arena[o:o + p.numel()].view_as(p),
                                 exp_avg_sq=self.v_arena[o:o + p.numel()].view_as(p))
            _OWNER[id(p)] = self
            o += p.numel()
        self.g_arena = None
        self._stats_valid = False
        self._scale = None          # GradScaler's device scalar while its unscale_ is pending in the Adam pass
        self._max_norm = 0.0
        self._ema = None
        self.ema_applied_for = -1

    # ---- gradient arena: where backward wrote the gradients (training.TrainPlan.flat_grad) ----------------------------
    def _grads(self) -> torch.Tensor:
        ps = self.param_groups[0]["params"]
        g0 = ps[0].grad
        if g0 is None:
            raise RuntimeError("FlatAdam.step: parameters have no gradients (call backward first)")
        if self.g_arena is None or self.g_arena.data_ptr() != g0.data_ptr():
            st = g0.untyped_storage()
            off = g0.storage_offset()
            if g0.dtype != torch.float32 or st.nbytes() // 4 - off < self.n:
                raise RuntimeError("FlatAdam: gradients are not views of one flat fp32 arena (the drop-in nets' backward "
                                   "writes them there; other models need torch.optim.Adam)")
            self.g_arena = torch.empty(0, dtype=torch.float32, device=self.device).set_(st, off, (self.n,))
        base = self.g_arena.data_ptr()
        for p, o in zip(ps, self.offsets):
            if p.grad is None or p.grad.data_ptr() != base + 4 * o:
                raise RuntimeError("FlatAdam: a gradient is missing or not at its offset of the flat gradient arena")
        return self.g_arena

    def _ensure_stats(self, scale: torch.Tensor | None = None):
        if self._stats_valid:
            return
        g = self._grads()
        self._scale = scale
        check(lib().srb_grad_stats(self.ctx.handle, g.data_ptr(), self.n, None if scale is None else scale.data_ptr(),
                                   self.stats.data_ptr(), self.work.data_ptr(), _stream_ptr(self.device)), "srb_grad_stats")
        self._stats_valid = True

    @property
    def found_inf(self) -> torch.Tensor:
        return self.stats[1:2]

    def total_norm(self) -> torch.Tensor:
        """2-norm of the (unscaled) gradients of this step: what clip_grad_norm_ returns."""
        self._ensure_stats()
        return self.stats[0].sqrt()

    def attach_ema(self, ema, arena: torch.Tensor):
        if arena.numel() != self.n:
            raise RuntimeError("FlatAdam.attach_ema: the EMA arena does not match the parameter arena")
        self._ema = (weakref.ref(ema), arena)

    @torch.no_grad()
    def step(self, closure=None):
        if closure is not None:
            raise NotImplementedError("FlatAdam.step: closures are not supported")
        self._ensure_stats()
        g = self.param_groups[0]
        ema_ptr, d = None, 0.0
        if self._ema is not None and self._ema[0]() is not None:
            e = self._ema[0]()
            d = float(e.decay(e.updates + 1))
            ema_ptr = self._ema[1].data_ptr()
            self.ema_applied_for = e.updates + 1
        check(lib().srb_adam_ema_step(self.ctx.handle, self.p_arena.data_ptr(), self.g_arena.data_ptr(), self.m_arena.data_ptr(),
                                      self.v_arena.data_ptr(), ema_ptr, self.n,
                                      None if self._scale is None else self._scale.data_ptr(), self.stats.data_ptr(),
                                      self.work.data_ptr(), float(self._max_norm), float(g["lr"]), float(g["betas"][0]), float(g["betas"][1]),
                                      float(g["eps"]), float(g["weight_decay"]), d, _stream_ptr(self.device)), "srb_adam_ema_step")
        self._stats_valid, self._scale, self._max_norm = False, None, 0.0
        from . import training
        training.PARAM_EPOCH[0] += 1            # the packed operands of the conv kernels are stale now
        return None

    def zero_grad(self, set_to_none: bool = True):
        """The gradients are views of the arena backward rewrites: dropping them is all there is to do.  set_to_none=False
        is treated the same way — a retained .grad would alias the arena the next backward writes (training.TrainPlan.backward
        refuses that), and gradient accumulation over several backward passes is not supported by the flat arena."""
        self._stats_valid, self._scale, self._max_norm = False, None, 0.0
        for p in self.param_groups[0]["params"]:
            p.grad = None

    def load_state_dict(self, state_dict):
        """torch's format; the loaded moments are copied INTO the flat arenas (the views must stay views)."""
        views = {id(p): (s["exp_avg"], s["exp_avg_sq"]) for p, s in self.state.items()}
        steps = self.stats[2]
        super().load_state_dict(state_dict)
        with torch.no_grad():
            if not self.state:             # a checkpoint taken before the first step: fresh moments, step 0
                self.m_arena.zero_(); self.v_arena.zero_(); steps.zero_()
            for p in self.param_groups[0]["params"]:
                m, v = views[id(p)]
                s = self.state.get(p)
                if s:
                    m.copy_(s["exp_avg"]); v.copy_(s["exp_avg_sq"])
                    steps.copy_(torch.as_tensor(s["step"], dtype=torch.float32))
                self.state[p] = dict(step=steps, exp_avg=m, exp_avg_sq=v)      # the views stay views of the arenas


def clip_grad_norm_(parameters, max_norm, norm_type=2.0, error_if_nonfinite=False, foreach=None):
    """torch.nn.utils.clip_grad_norm_ (train.py:59).  For the parameters of a FlatAdam: returns the global norm (device
    scalar, no sync) and defers the scaling to the optimizer's pass over the gradients."""
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    ps = list(parameters)
    opt = _OWNER.get(id(ps[0])) if ps else None
    if (opt is None or float(norm_type) != 2.0 or error_if_nonfinite or len(ps) != len(opt.offsets)
            or any(_OWNER.get(id(p)) is not opt for p in ps)):
        return torch.nn.utils.clip_grad_norm_(ps, max_norm, norm_type=norm_type, error_if_nonfinite=error_if_nonfinite, foreach=foreach)
    norm = opt.total_norm()
    opt._max_norm = float(max_norm)
    return norm


class GradScaler(torch.amp.GradScaler):
    """torch.amp.GradScaler (train.py:12,57-61) that hands a FlatAdam its scale instead of rewriting the gradients."""

    def __init__(self, device="cuda", **kw):
        super().__init__(device, **kw)

    def unscale_(self, optimizer):
        if not self._enabled or not isinstance(optimizer, FlatAdam):
            return super().unscale_(optimizer)
        self._check_scale_growth_tracker("unscale_")
        st = self._per_optimizer_states[id(optimizer)]
        if st["stage"] is OptState.UNSCALED:
            raise RuntimeError("unscale_() has already been called on this optimizer since the last update().")
        if st["stage"] is OptState.STEPPED:
            raise RuntimeError("unscale_() is being called after step().")
        optimizer._stats_valid = False
        optimizer._ensure_stats(self._scale)
        st["found_inf_per_device"] = {optimizer.device: optimizer.found_inf}
        st["stage"] = OptState.UNSCALED

    def step(self, optimizer, *args, **kwargs):
        if not self._enabled or not isinstance(optimizer, FlatAdam):
            return super().step(optimizer, *args, **kwargs)
        self._check_scale_growth_tracker("step")
        st = self._per_optimizer_states[id(optimizer)]
        if st["stage"] is OptState.STEPPED:
            raise RuntimeError("step() has already been called since the last update().")
        if st["stage"] is OptState.READY:
            self.unscale_(optimizer)
        ret = optimizer.step(*args, **kwargs)      # skipped on the device when found_inf is set
        st["stage"] = OptState.STEPPED
        return ret
