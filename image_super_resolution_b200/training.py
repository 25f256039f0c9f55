"""Training step of the class nets on the sm_100a kernels: forward with saved activations, backward (data, weight and
bias gradients), wrapped as one torch.autograd.Function so that the reference's `train()` (train.py:41-67: autocast
forward, loss, GradScaler backward, clip, Adam, EMA) runs unmodified on the drop-in `EResNet`.

Schedule, all launches through the C ABI:
  forward   the inference schedule of models._NetPlan, but every RDB owns its 192-channel dense buffer (saved for
            backward; torch.cat stays free) and the operands are 16-bit re-packed on the device from the fp32 masters.
            Operand type: fp16 — the reference's own training precision (train.py:54 autocast + GradScaler).
            Measured on EResNet(16) (tools/gpu_train_parity.py): with GradScaler's loss scale (65536) every one of the
            488 gradient tensors has cosine >= 0.9991 vs fp32 autograd; bf16 operands reach only 0.9958 (242 tensors
            below the 0.999 contract), fp16 WITHOUT loss scaling underflows.  `net.train_operand_dtype = "bf16"` selects bf16;
  backward  per conv: weight gradient (tcgen05: srb_grad_batch_run, one launch for all convs; the 9x9 head/tail in the same mapping, KS = 9),
            bias gradient, and the data gradient as a conv of dY with the transposed / rotated filter that
            ACCUMULATES in place into the dense block's gradient buffer and applies the LeakyReLU mask of the slice
            that becomes final with that launch (no separate elementwise kernels inside the trunk).
  DDP       the flat fp32 gradient arena is all-reduced once (NCCL over NVLink) before autograd hands the gradients
            to the optimizer, i.e. between train.py:57 and :58.
There is no PyTorch / cuDNN fallback: anything unsupported raises.
"""
from __future__ import annotations

import torch
import torch.distributed as dist
from torch import nn

from . import ops, _lib
from ._lib import Context


PARAM_EPOCH = [0]      # bumped by optimizers that update the parameters behind autograd's version counters (optim.FlatAdam)


def _chunks(c: int) -> list[int]:
    """Output-channel chunks of a data-gradient conv with c output channels: one launch up to 160 channels (N = 96 .. 160
    MMAs run at the full tensor rate, N <= 64 ones at 40-67 %), two 96-channel halves for the 192-channel gradient of an
    RDB's last conv (its 64 -> 192 filter does not fit in shared memory in one piece)."""
    return [c] if c <= 160 else [96] * (c // 96) + ([c % 96] if c % 96 else [])


def _slope_of(act) -> float:
    if isinstance(act, nn.LeakyReLU):
        return float(act.negative_slope)
    if isinstance(act, (nn.Identity, nn.Tanh)):
        return 1.0
    raise NotImplementedError(f"activation {act.__class__.__name__} has no fused backward")


class TrainPlan:
    """Device-packed operands, saved activations and the fwd/bwd launch schedule of one EResNet on one device."""

    allow_bn = False           # BNTrainPlan (below) schedules the nets whose `Conv` blocks carry a BatchNorm2d

    def __init__(self, net, device: torch.device, dtype="fp16", allreduce: bool = True):
        if not self.allow_bn and any(isinstance(m, nn.BatchNorm2d) for m in net.modules()):
            raise RuntimeError("TrainPlan schedules the BatchNorm-free EResNet; nets with BatchNorm2d train through BNTrainPlan")
        self.net, self.device, self.dt = net, device, ops.dtype_code(dtype)
        self.ctx = Context.get(device.index)
        self.allreduce = allreduce
        mk = lambda co, ci, k: ops.DeviceConv(self.ctx, co, ci, k, self.dt)  # noqa: E731
        # 1: nine-tap kernels (default: 8.73 ms per step at B = 16 x 64 x 64); 2: dy taps folded into N (csrc/srb_conv.cuh MODE_DYF: 9.12 ms —
        # fewer, larger MMAs but 14-row tiles and a heavier epilogue; ahead only on large frames). profiles/chain_variants_r2.log
        fold = int(getattr(net, "train_conv_fold", 1))
        mkw = lambda co, ci, k: ops.DeviceConv(self.ctx, co, ci, k, self.dt, ws=fold)  # noqa: E731  (layers of a ConvChain)
        with torch.cuda.device(device):
            self.f_conv0 = mk(64, 3, 9)
            self.rdbs = []
            for rrdb in net.residual:
                for rdb in rrdb.net:
                    fwd = [mkw(32, 64 + 32 * i, 3) for i in range(4)] + [mk(64, 192, 3)]
                    # data gradients, gathered: stage k = ONE conv over the concatenated output gradients of all later convs
                    # [dY4 (64) | dY3 | .. | dY_{k+1}] -> the 32 channels of o_k; fin = all five -> the 64 input channels
                    stage = [mkw(32, 64 + 32 * (3 - k), 3) for k in range(4)]
                    fin = mk(64, 192, 3)
                    self.rdbs.append((rrdb, rdb, fwd, stage, fin))
            self.f_conv1, self.d_conv1 = mk(64, 64, 3), mk(64, 64, 3)
            self.f_scaler = [[mk(64, 64, 3) for _ in range(4)] for _ in net.scaler]
            self.d_scaler = [mk(64, 256, 3) for _ in net.scaler]
            self.f_conv2, self.d_conv2 = mk(3, 64, 9), mk(64, 3, 9)
        self.params = list(net.parameters())
        self.numel = sum(p.numel() for p in self.params)
        # one flat arena (zero-padded to whole 16-byte granules): the all-reduce, clip_grad_norm_ and optim.FlatAdam see it whole
        self.flat_grad = torch.zeros((self.numel + 3) // 4 * 4, dtype=torch.float32, device=device)
        self.grad_views, o = {}, 0
        for p in self.params:
            self.grad_views[id(p)] = self.flat_grad[o:o + p.numel()].view_as(p)
            o += p.numel()
        self.versions = None
        self.buf = {}
        self.saved = None
        self.use_graphs = True
        self.use_pdl = True          # programmatic edges between the conv nodes of the captured graphs
        # conv0 .. conv3 of an RDB (and the four gathered data-gradient stages of its backward) as ONE launch each: persistent
        # CTAs with tile-level dependencies instead of four launches (ops.ConvChain; csrc/srb_conv.cuh CHAIN)
        self.use_chains = True
        self.fchains, self.bchains, self.chain_key = None, None, None
        # the block's LAST conv (192 -> 64 with its residual terms; in the backward the data gradient of the block's input) rides in the
        # chain as a fifth layer on the 64-output chain kernel (nine-tap layout only)
        self.chain5 = bool(getattr(net, "train_chain5", 1)) and fold == 1
        self.chain_nt4 = bool(getattr(net, "train_chain_nt4", False))    # super-tiles of four M-tiles in the chains even for small batches
        self.wgrad_splitk = 1
        # Internal gradient shift: the data-gradient chain runs on 16-bit tensors whose magnitudes span the tail conv (~1 with
        # GradScaler's 65536) down to the dense blocks of a x4 net at initialisation (~4e-5: fp16 subnormals).  The chain is
        # linear, so the upstream gradient is multiplied by 2^8 where it enters (tanh backward) and every weight / bias
        # gradient by 2^-8 where it leaves (exact: powers of two).  An overflow shows up as inf in the fp32 gradients, which
        # GradScaler answers by skipping the step and halving its scale, exactly as for its own scale.
        self.grad_shift = 256.0
        self.graphs = {}
        self.pack_batch, self.pack_ptrs = None, None
        self.grad_batch, self.grad_key, self._gjobs = None, None, None
        # Data parallel, optional (ddp_parts > 1): the deferred weight-gradient batch runs as `ddp_parts` launches over contiguous
        # slices of the gradient arena (last slice first: backpropagation finishes the net's tail first), and the all-reduce
        # of a finished slice is handed to NCCL (async, its own stream) while the next launch runs.  MEASURED on 8 B200
        # (profiles/ddp_overlap_r2.log): 2 parts 10.62 ms per step against 9.84 ms with ONE all-reduce after the whole backward
        # (9.55 ms on one GPU) — the NCCL kernel and the persistent weight-gradient CTAs (one per SM, 226 KB of shared memory)
        # slow each other down by more than the 0.23 ms the all-reduce costs on its own, and two one-wave launches pack the
        # SMs worse than one 2.7-wave launch.  Default: one all-reduce.
        self.ddp_parts = 1
        self.grad_parts = None          # [(GradBatch, arena lo, arena hi)] in launch order

    # ---- operands -------------------------------------------------------------------------------------------
    def _versions(self):
        return tuple(p._version for p in self.params) + (PARAM_EPOCH[0],)

    def __deepcopy__(self, memo):
        return None            # deepcopy(model) (ModelEMA, utils/models.py:25) must not clone device handles; a copy re-plans

    def _pack_jobs(self):
        n = self.net
        bias = lambda m: None if m.conv.bias is None else m.conv.bias.data     # noqa: E731  (`Conv` blocks are bias-free)
        jobs = [(self.f_conv0, n.conv0.conv.weight.data, n.conv0.conv.bias.data, 0, 0, False)]
        for r, (rrdb, rdb, fwd, stage, fin) in enumerate(self.rdbs):
            Wt = []
            for i in range(5):
                m = getattr(rdb, f"conv{i}") if i < 4 else rdb.conv
                w, b = m.conv.weight.data, bias(m)
                jobs.append((fwd[i], w, b, 0, 0, False))
                Wt.append(w)
            # transposed / rotated filters of the gathered data-gradient convs: input channels [0,64) come from conv4 (times
            # the block's output scale), [64 + 32(3-j), +32) from conv_j; output channels = source input channels o_off..
            s_out = float(rdb.add_rate) * (float(rrdb.add_rate) if r % 3 == 2 else 1.0)
            for k in range(4):
                o_off = 64 + 32 * k
                jobs.append((stage[k], Wt[4], None, o_off, 0, True, 0, 64, s_out))
                for j in range(3, k, -1):
                    jobs.append((stage[k], Wt[j], None, o_off, 0, True, 64 + 32 * (3 - j), 32, 1.0))
            jobs.append((fin, Wt[4], None, 0, 0, True, 0, 64, s_out))
            for j in range(3, -1, -1):
                jobs.append((fin, Wt[j], None, 0, 0, True, 64 + 32 * (3 - j), 32, 1.0))
        jobs.append((self.f_conv1, n.conv1.conv.weight.data, bias(n.conv1), 0, 0, False))
        jobs.append((self.d_conv1, n.conv1.conv.weight.data, None, 0, 0, True))
        for s, fq, dq in zip(n.scaler, self.f_scaler, self.d_scaler):
            w, b = s.net[0].conv.weight.data, s.net[0].conv.bias.data
            for q in range(4):
                jobs.append((fq[q], w, b, 64 * q, 0, False))
            jobs.append((dq, w, None, 0, 0, True))
        jobs.append((self.f_conv2, n.conv2.conv.weight.data, n.conv2.conv.bias.data, 0, 0, False))
        jobs.append((self.d_conv2, n.conv2.conv.weight.data, None, 0, 0, True))
        return jobs

    def repack(self, ptrs=None, track_versions=True):
        """fp32 master weights (device) -> 16-bit UMMA operands, forward and data-gradient layouts: one launch."""
        if ptrs is None:
            ptrs = tuple(p.data_ptr() for p in self.params)
        if self.pack_batch is None or self.pack_ptrs != ptrs:
            for p in self.params:
                if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                    raise RuntimeError("training needs contiguous fp32 CUDA parameters (model.to(device) first)")
            self.pack_batch, self.pack_ptrs = ops.PackBatch(self.ctx, self._pack_jobs()), ptrs
        self.pack_batch.run(self.device)
        if track_versions:             # (eager mode re-packs only when a parameter changed; graph mode re-packs every forward)
            self.versions = self._versions()

    def _mk_chain(self, layers, keep=()):
        if self.chain_nt4:
            self.ctx.set_option(_lib.OPT_PATH, _lib.PATH_LARGE_TILING)
        try:
            return ops.ConvChain(self.ctx, layers, keep=keep)
        finally:
            if self.chain_nt4:
                self.ctx.set_option(_lib.OPT_PATH, 0)

    def _b(self, name, N, planes, H, W):
        key = (name, N, planes, H, W)
        t = self.buf.get(name)
        if t is None or t[0] != key:
            t = (key, ops.c8_empty(N, planes, H, W, self.dt, self.device))
            self.buf[name] = t
        return t[1]

    def _pdl(self, on: bool):
        if self.use_pdl:
            _lib.check(_lib.lib().srb_set_programmatic_launch(self.ctx.handle, 1 if on else 0), "srb_set_programmatic_launch")

    # ---- CUDA graphs: the step is ~2 000 launches of 5-20 us kernels, host-bound from Python otherwise ------------
    def _graph_key(self, x):
        return (tuple(x.shape), tuple(p.data_ptr() for p in self.params))      # one pass over the parameters per forward

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not self.use_graphs:
            if self.versions != self._versions():
                self.repack()
            return self._forward_impl(x)
        key = self._graph_key(x)
        st = self.graphs.get("key")
        if st != key:                                        # new shape / re-allocated parameters: capture again
            self.graphs = {"key": key, "x": x.clone()}
            self.repack()
            keep = self._warmup_state()
            self._forward_impl(self.graphs["x"])             # eager warm-up: allocates buffers, sets kernel attributes
            torch.cuda.synchronize(self.device)
            self._warmup_restore(keep)
            # The operand re-pack stays outside the graph: inside it the packed weights are then constants, and the conv
            # nodes can be linked by programmatic edges (each conv's prologue overlaps its predecessor's tail).
            g = torch.cuda.CUDAGraph()
            self._pdl(True)
            try:
                with torch.cuda.graph(g, capture_error_mode="thread_local"):
                    self.graphs["out"] = self._forward_impl(self.graphs["x"])
            finally:
                self._pdl(False)
            self.graphs["fwd"] = g
            self.graphs["saved"] = self.saved
        self.repack(key[1], track_versions=False)     # the host path between two steps is GPU idle time (loss.item() syncs)
        self.graphs["x"].copy_(x)
        self.graphs["fwd"].replay()
        self.saved = self.graphs["saved"]
        return self.graphs["out"]

    def _warmup_state(self):
        """State the eager warm-up forward must not change (none here; BatchNorm running statistics in BNTrainPlan)."""
        return None

    def _warmup_restore(self, keep):
        pass

    def backward(self, gout: torch.Tensor) -> list[torch.Tensor]:
        # The gradients handed to autograd are views of the flat arena this call rewrites.  A parameter that still holds
        # last step's .grad holds the SAME memory: AccumulateGrad would add the new gradient onto itself (2 g instead of
        # g1 + g2).  Gradient accumulation over several backward passes is not supported: refuse instead of being wrong.
        for p in self.params:
            if p.grad is not None:
                raise RuntimeError("backward with .grad still set: the drop-in nets write their gradients into one flat arena, "
                                   "so gradients cannot be accumulated over several backward passes — call "
                                   "optimizer.zero_grad() (set_to_none=True, torch's default and what train.py:53 does) first")
        if not self.use_graphs:
            grads = self._backward_impl(gout)
        else:
            if self.saved is None:
                raise RuntimeError("backward without a saved forward")
            if "bwd" not in self.graphs:
                self.graphs["gout"] = gout.float().contiguous().clone()
                self._backward_impl(self.graphs["gout"], run_batch=False)     # (also builds the gradient batch / its parts)
                torch.cuda.synchronize(self.device)
                g = torch.cuda.CUDAGraph()
                self._pdl(True)
                try:
                    with torch.cuda.graph(g, capture_error_mode="thread_local"):
                        self.graphs["grads"] = self._backward_impl(self.graphs["gout"], run_batch=False)
                finally:
                    self._pdl(False)
                self.graphs["bwd"] = g
            self.graphs["gout"].copy_(gout)
            self.graphs["bwd"].replay()
            grads = self.graphs["grads"]
            if self.grad_parts is not None:      # data parallel: weight-gradient launches outside the graph, all-reduce overlapped
                self._run_parts()
        self.saved = None
        if self.grad_parts is None and self._ddp():
            allreduce_mean_(self.flat_grad)
        return grads

    # ---- forward --------------------------------------------------------------------------------------------
    def _forward_impl(self, x: torch.Tensor) -> torch.Tensor:
        n = self.net
        N, _, H, W = x.shape
        cv = ops.conv2d_ex
        xin = self._b("xin", N, 1, H, W)
        ops.nchw_to_c8(x, out=xin, planes=1, dtype=self.dt)
        nr = len(self.rdbs)
        dense = [self._b(f"dense{r}", N, 24, H, W) for r in range(nr)]
        tr = self._b("tr", N, 8, H, W)
        s0 = _slope_of(n.conv0.act)
        cv(self.f_conv0, xin, 1, 0, dense[0], 24, 0, N, H, W, slope=s0, flags=ops.F_LRELU)     # x0 = planes 0..7 of dense[0]
        def last_conv_kw(r):
            """residual terms of an RDB's last conv: add_rate*conv + rdb_in (utils/models.py:271), for the third RDB of an RRDB
            R*(add_rate*conv + rdb_in) + rrdb_in (:317)"""
            rrdb, rdb = self.rdbs[r][0], self.rdbs[r][1]
            a = float(rdb.add_rate)
            if r % 3 < 2:
                return dict(residual=dense[r], res_planes=24, res_off=0, res_scale=1.0, out_scale=a)
            R = float(rrdb.add_rate)
            return dict(residual=dense[r], res_planes=24, res_off=0, res_scale=R, out_scale=R * a, res2=dense[r - 2], res2_planes=24,
                        res2_off=0, res2_scale=1.0)
        if self.use_chains and self.chain_key != (N, H, W):
            self.fchains, self.bchains, self.chain_key = [], None, (N, H, W)
            for r, (rrdb, rdb, fwd, _, _) in enumerate(self.rdbs):
                D = dense[r]
                dst, dplanes = (dense[r + 1], 24) if r + 1 < nr else (tr, 8)
                layers = [(fwd[i], ops.conv_args(D, 24, 0, D, 24, 8 + 4 * i, N, H, W, slope=_slope_of(getattr(rdb, f"conv{i}").act),
                                                 flags=ops.F_LRELU)) for i in range(4)]
                if self.chain5:     # ... and the block's last conv (64 outputs, residual terms) as the chain's fifth layer
                    layers.append((fwd[4], ops.conv_args(D, 24, 0, dst, dplanes, 0, N, H, W, **last_conv_kw(r))))
                self.fchains.append(self._mk_chain(layers, keep=(D, dst, dense[r - 2] if r % 3 == 2 else None)))
        for r, (rrdb, rdb, fwd, _, _) in enumerate(self.rdbs):
            D = dense[r]
            dst, dplanes = (dense[r + 1], 24) if r + 1 < nr else (tr, 8)
            if self.use_chains:
                self.fchains[r].run(self.device)
                if self.chain5:
                    continue
            else:
                for i in range(4):
                    si = _slope_of(getattr(rdb, f"conv{i}").act)
                    cv(fwd[i], D, 24, 0, D, 24, 8 + 4 * i, N, H, W, slope=si, flags=ops.F_LRELU)
            a = float(rdb.add_rate)
            if r % 3 < 2:    # add_rate*conv + rdb_in  (utils/models.py:271)
                cv(fwd[4], D, 24, 0, dst, dplanes, 0, N, H, W, residual=D, res_planes=24, res_off=0, res_scale=1.0, out_scale=a)
            else:            # R*(add_rate*conv + rdb_in) + rrdb_in  (utils/models.py:317)
                R = float(rrdb.add_rate)
                cv(fwd[4], D, 24, 0, dst, dplanes, 0, N, H, W, residual=D, res_planes=24, res_off=0, res_scale=R,
                   out_scale=R * a, res2=dense[r - 2], res2_planes=24, res2_off=0, res2_scale=1.0)
        t = self._b("t", N, 8, H, W)
        cv(self.f_conv1, tr, 8, 0, t, 8, 0, N, H, W, residual=dense[0], res_planes=24, res_off=0)   # + long skip x0
        cur, h, w, ups = t, H, W, []
        for k, (s, fq) in enumerate(zip(n.scaler, self.f_scaler)):
            up = self._b(f"up{k}", N, 8, 2 * h, 2 * w)
            sl = _slope_of(s.net[2])
            for q in range(4):
                cv(fq[q], cur, 8, 0, up, 8, 2 * q, N, h, w, slope=sl, flags=ops.F_LRELU | ops.F_PIXSHUF2)
            ups.append((cur, up, h, w))
            cur, h, w = up, 2 * h, 2 * w
        out = torch.empty((N, 3, h, w), dtype=torch.float32, device=self.device)
        tanh = ops.F_TANH if isinstance(n.conv2.act, nn.Tanh) else 0
        cv(self.f_conv2, cur, 8, 0, out, 0, 0, N, h, w, flags=tanh | ops.F_F32OUT)
        self.saved = dict(N=N, H=H, W=W, xin=xin, dense=dense, tr=tr, t=t, ups=ups, out=out, tanh=bool(tanh))
        return out

    # ---- backward -------------------------------------------------------------------------------------------
    def _wb(self, m, x, x_planes, x_off, cin, dy, dy_planes, dy_off, cout, N, H, W, scale=1.0):
        """weight + bias gradient of a 3x3 conv module: recorded once as a job of the deferred batch (run at the end of
        backward: weight gradients are off the critical path and every unit is then owned by one CTA)."""
        if self._gjobs is not None:
            self._gjobs.append(dict(x=x, x_planes=x_planes, x_off=x_off, cin=cin, dy=dy, dy_planes=dy_planes, dy_off=dy_off,
                                    cout=cout, dw=self.grad_views[id(m.conv.weight)],
                                    db=None if m.conv.bias is None else self.grad_views[id(m.conv.bias)],
                                    scale=scale / float(self.grad_shift), N=N, H=H, W=W))

    def _backward_impl(self, gout: torch.Tensor, run_batch: bool = True) -> list[torch.Tensor]:
        sv, n = self.saved, self.net
        if sv is None:
            raise RuntimeError("backward without a saved forward")
        N, H, W = sv["N"], sv["H"], sv["W"]
        cv = ops.conv2d_ex
        self.flat_grad.zero_()
        gkey = (N, H, W, tuple(sv["out"].shape))
        self._gjobs = [] if (self.grad_batch is None or self.grad_key != gkey) else None
        gout = gout.float().contiguous()
        h2, w2 = sv["out"].shape[2], sv["out"].shape[3]
        # ---- tail conv: tanh', 9x9 weight gradient, data gradient = 9x9 conv 3 -> 64 (head kernel, transposed filter)
        if not sv["tanh"]:
            raise NotImplementedError("tail conv without tanh")
        sh, ish = float(self.grad_shift), 1.0 / float(self.grad_shift)
        g2 = ops.tanh_bwd_to_c8(gout, sv["out"], self.dt, scale=sh, g8=self._b("g2", N, 1, h2, w2))
        up_last = sv["ups"][-1][1] if sv["ups"] else sv["t"]
        ops.conv9_wgrad(up_last, 8, 0, 64, g2, 1, 0, 3, self.grad_views[id(n.conv2.conv.weight)], N, h2, w2, scale=ish)
        if self._gjobs is not None:
            self._gjobs.append(dict(dy=g2, dy_planes=1, dy_off=0, cout=3, db=self.grad_views[id(n.conv2.conv.bias)], N=N, H=h2, W=w2,
                                    scale=ish))
        g_cur = self._b("g_up_last", N, 8, h2, w2)
        cv(self.d_conv2, g2, 1, 0, g_cur, 8, 0, N, h2, w2)
        # ---- scalers, last to first: LeakyReLU' + PixelShuffle inverse, wgrad 64 -> 256, dgrad 256 -> 64
        for k in range(len(sv["ups"]) - 1, -1, -1):
            cur_in, up, h, w = sv["ups"][k]
            s = n.scaler[k]
            g_s = self._b(f"g_s{k}", N, 32, h, w)
            ops.pixunshuffle_mask(g_cur, up, g_s, _slope_of(s.net[2]))
            self._wb(s.net[0], cur_in, 8, 0, 64, g_s, 32, 0, 256, N, h, w)
            g_in = self._b(f"g_in{k}", N, 8, h, w)
            cv(self.d_scaler[k], g_s, 32, 0, g_in, 8, 0, N, h, w)
            g_cur = g_in
        g_t = g_cur                                         # gradient of t = x0 + conv1(trunk)
        # ---- conv1
        self._wb(n.conv1, sv["tr"], 8, 0, 64, g_t, 8, 0, 64, N, H, W)
        # ---- RRDBs, last to first.  Every RDB owns a dense GRADIENT buffer G (24 planes, kept: the deferred weight gradients
        # read it) that mirrors the forward's dense activation buffer: planes 0..7 = gy (gradient of the RDB's output, written
        # by the block behind it), then dY3, dY2, dY1, dY0 (gradients of the pre-activations of o3 .. o0), each produced by
        # one conv over everything to its left — no read-modify-write passes, fp32 accumulation over all sources.
        nr = len(self.rdbs)
        Gd = [self._b(f"G{j}", N, 24, H, W) for j in range(nr)]
        gxf = self._b("g_x0t", N, 8, H, W)                  # gradient of x0 through the trunk
        if self.use_chains and self.bchains is None:
            self.bchains = []
            for r in range(nr):
                rrdb, rdb, _, stage, _ = self.rdbs[r]
                D, G = sv["dense"][r], Gd[r]
                layers = [(stage[k], ops.conv_args(G, 24, 0, G, 24, 8 + 4 * (3 - k), N, H, W, mask=D, mask_planes=24, mask_off=8 + 4 * k,
                                                   mask_p0=0, mask_p1=4, mask_slope=_slope_of(getattr(rdb, f"conv{k}").act)))
                          for k in range(3, -1, -1)]
                if self.chain5:     # ... and the data gradient of the block's input (all five sources -> 64 channels, + s_res * gy [+ RRDB skip])
                    fin = self.rdbs[r][4]
                    dstb, dpl = (Gd[r - 1], 24) if r > 0 else (gxf, 8)
                    s_res = float(rrdb.add_rate) if r % 3 == 2 else 1.0
                    kwb = dict(residual=G, res_planes=24, res_off=0, res_scale=s_res)
                    if r % 3 == 0:
                        kwb.update(res2=Gd[r + 2], res2_planes=24, res2_off=0, res2_scale=1.0)
                    layers.append((fin, ops.conv_args(G, 24, 0, dstb, dpl, 0, N, H, W, **kwb)))
                self.bchains.append(self._mk_chain(layers, keep=(D, G, Gd[r - 1] if r > 0 else gxf, Gd[r + 2] if r % 3 == 0 else None)))
        cv(self.d_conv1, g_t, 8, 0, Gd[nr - 1], 24, 0, N, H, W)
        for r in range(nr - 1, -1, -1):
            rrdb, rdb, _, stage, fin = self.rdbs[r]
            D, G = sv["dense"][r], Gd[r]
            d = r % 3
            a, R = float(rdb.add_rate), float(rrdb.add_rate)
            s_out, s_res = (a * R, R) if d == 2 else (a, 1.0)          # utils/models.py:271 / :317
            sl = [_slope_of(getattr(rdb, f"conv{i}").act) for i in range(4)]
            self._wb(rdb.conv, D, 24, 0, 192, G, 24, 0, 64, N, H, W, scale=s_out)
            if self.use_chains:
                self.bchains[r].run(self.device)
            for k in range(3, -1, -1):
                nin = 8 + 4 * (3 - k)                                   # gy + dY3 .. dY_{k+1}
                if not self.use_chains:
                    cv(stage[k], G, 24, 0, G, 24, nin, N, H, W, mask=D, mask_planes=24, mask_off=8 + 4 * k, mask_p0=0, mask_p1=4,
                       mask_slope=sl[k])
                self._wb(getattr(rdb, f"conv{k}"), D, 24, 0, 64 + 32 * k, G, 24, nin, 32, N, H, W)
            if self.use_chains and self.chain5:
                continue                                                # (fin is the chain's fifth layer)
            dst, dplanes = (Gd[r - 1], 24) if r > 0 else (gxf, 8)
            kw = dict(residual=G, res_planes=24, res_off=0, res_scale=s_res)
            if d == 0:                                                  # + the RRDB skip (utils/models.py:317)
                kw.update(res2=Gd[r + 2], res2_planes=24, res2_off=0, res2_scale=1.0)
            cv(fin, G, 24, 0, dst, dplanes, 0, N, H, W, **kw)
        # ---- head conv: total gradient of x0 = trunk path + long skip, LeakyReLU', 9x9 weight gradient
        gx0 = self._b("gx0", N, 8, H, W)
        ops.c8_axpby_mask(gxf, 8, 0, 1.0, g_t, 8, 0, 1.0, sv["dense"][0], 24, 0, _slope_of(n.conv0.act), gx0, 8, 0, 8, N, H, W)
        ops.conv9_wgrad(sv["xin"], 1, 0, 3, gx0, 8, 0, 64, self.grad_views[id(n.conv0.conv.weight)], N, H, W, scale=ish)
        if self._gjobs is not None:
            self._gjobs.append(dict(dy=gx0, dy_planes=8, dy_off=0, cout=64, db=self.grad_views[id(n.conv0.conv.bias)], N=N, H=H, W=W,
                                    scale=ish))
            self.grad_batch, self.grad_key = ops.GradBatch(self.ctx, self._gjobs, self.dt), gkey
            self.grad_parts = self._split_jobs(self._gjobs) if self._ddp() and self.ddp_parts > 1 else None
            self._gjobs = None
        if self.grad_parts is None:
            self.grad_batch.run(self.device, self.wgrad_splitk)
        elif run_batch:
            self._run_parts()
        return [self.grad_views[id(p)] for p in self.params]

    def _ddp(self) -> bool:
        return self.allreduce and dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1

    def _split_jobs(self, jobs):
        """Partition the weight / bias gradient jobs by the arena offset of their targets into `ddp_parts` contiguous slices of
        about equal work (FLOPs of the weight-gradient GEMMs); returned in launch order = highest arena slice first."""
        base, esz = self.flat_grad.data_ptr(), 4

        def off(j):
            t = j.get("dw") if j.get("dw") is not None else j.get("db")
            return (t.data_ptr() - base) // esz

        def cost(j):
            return j["N"] * j["H"] * j["W"] * j.get("cin", 0) * j["cout"] + 1

        js = sorted(jobs, key=off)
        total, parts, cur, acc = sum(cost(j) for j in js), [], [], 0
        for j in js:
            cur.append(j)
            acc += cost(j)
            if len(parts) < self.ddp_parts - 1 and acc >= total * (len(parts) + 1) / self.ddp_parts:
                parts.append(cur)
                cur = []
        if cur:
            parts.append(cur)
        # slice boundaries: a slice ends where the next part's first target begins (parameters without a deferred job — the 9x9
        # filters, computed inside the captured chain — fall into the slice that surrounds them)
        bounds = [0] + [off(p[0]) for p in parts[1:]] + [self.flat_grad.numel()]
        out = [(ops.GradBatch(self.ctx, p, self.dt), bounds[i], bounds[i + 1]) for i, p in enumerate(parts)]
        return out[::-1]

    def _run_parts(self):
        """The weight-gradient launches with the all-reduce of every finished arena slice overlapping the next launch."""
        works = []
        for gb, lo, hi in self.grad_parts:
            gb.run(self.device, self.wgrad_splitk)
            works.append(allreduce_mean_(self.flat_grad[lo:hi], async_op=True))
        for w in works:
            if w is not None:
                w.wait()
        if dist.get_backend() != "nccl":
            self.flat_grad.div_(dist.get_world_size())


def allreduce_mean_(flat: torch.Tensor, async_op: bool = False):
    """All-reduce of (a slice of) the flat gradient arena (NCCL over NVLink); mean over ranks like DDP.
    async_op: returns the work handle (NCCL: AVG, overlaps the caller's stream; gloo: SUM — the caller divides afterwards)."""
    ws = dist.get_world_size()
    if dist.get_backend() == "nccl":
        w = dist.all_reduce(flat, op=dist.ReduceOp.AVG, async_op=async_op)
        return w if async_op else flat
    w = dist.all_reduce(flat, op=dist.ReduceOp.SUM, async_op=async_op)
    if async_op:
        return w
    flat.div_(ws)
    return flat


class BNTrainPlan(TrainPlan):
    """Training step of `ResNet` (utils/models.py:592-618): every conv of the trunk is a `Conv` block, conv -> BatchNorm2d -> act
    (:75-111), in training mode — batch statistics, running-statistics update, and the BatchNorm backward.

    A train-mode BatchNorm needs the whole batch before its first output, so neither the activation (forward) nor the
    LeakyReLU mask (backward) can ride in the conv's epilogue, and the tile-level conv chains do not apply: every `Conv` is
        forward   conv (raw y, saved) -> srb_bn_stats -> srb_bn_apply (normalise, act, and the block's residual terms)
        backward  gathered data-gradient conv (raw dL/da) -> srb_bn_bwd_reduce -> srb_bn_bwd_apply (dL/dy, d gamma, d beta)
    on the same conv kernels, packed operands, dense activation / gradient buffers, deferred weight-gradient batch, flat
    gradient arena and CUDA graphs as TrainPlan.  Per-rank statistics under data parallelism (what torch DDP does without
    SyncBatchNorm; the reference has no multi-GPU training at all)."""
    allow_bn = True

    def __init__(self, net, device, dtype="fp16", allreduce: bool = True):
        super().__init__(net, device, dtype, allreduce)
        self.use_chains = False
        self.bn_modules = [m for m in net.modules() if isinstance(m, nn.BatchNorm2d)]
        for m in self.bn_modules:
            if not m.affine or not m.track_running_stats:
                raise NotImplementedError("BatchNorm2d without affine parameters / running statistics")
        nbn = 5 * len(self.rdbs) + 1
        self.bn_sums = torch.zeros((nbn, 2, 64), dtype=torch.float64, device=device)      # sum y, sum y^2 per `Conv`
        self.bn_bsums = torch.zeros((nbn, 2, 64), dtype=torch.float64, device=device)     # sum dz, sum dz * x_hat
        self.bn_pass, self.bn_key = None, None

    def _warmup_state(self):
        return [(m, m.running_mean.clone(), m.running_var.clone(), m.num_batches_tracked.clone()) for m in self.bn_modules]

    def _warmup_restore(self, keep):
        for m, rm, rv, nb in keep:
            m.running_mean.copy_(rm); m.running_var.copy_(rv); m.num_batches_tracked.copy_(nb)

    def _mk_bn(self, idx, m, N, H, W, planes, y, y_planes, y_off, **kw):
        act = 1 if isinstance(m.act, nn.LeakyReLU) else 0
        if not isinstance(m.act, (nn.LeakyReLU, nn.Identity)):
            raise NotImplementedError(f"activation {m.act.__class__.__name__} after BatchNorm")
        return ops.BatchNormPass(self.ctx, self.dt, N, H, W, planes, y, y_planes, y_off, m.bn.weight.data, m.bn.bias.data,
                                 self.bn_sums[idx], self.bn_bsums[idx], bn=m.bn, slope=_slope_of(m.act) if act else 0.0, act=act, **kw)

    def _forward_impl(self, x: torch.Tensor) -> torch.Tensor:
        n = self.net
        N, _, H, W = x.shape
        cv = ops.conv2d_ex
        PARAM_EPOCH[0] += 1            # the running statistics move: inference plans derived from them are stale
        xin = self._b("xin", N, 1, H, W)
        ops.nchw_to_c8(x, out=xin, planes=1, dtype=self.dt)
        nr = len(self.rdbs)
        dense = [self._b(f"dense{r}", N, 24, H, W) for r in range(nr)]
        raw = [self._b(f"raw{r}", N, 24, H, W) for r in range(nr)]      # y of conv4 in planes 0..7, of conv_i in planes 8+4i..
        tr, t, rawc1 = self._b("tr", N, 8, H, W), self._b("t", N, 8, H, W), self._b("rawc1", N, 8, H, W)
        if self.bn_key != (N, H, W):
            self.bn_pass, self.bn_key = [], (N, H, W)
            for r, (rrdb, rdb, _, _, _) in enumerate(self.rdbs):
                D, Y = dense[r], raw[r]
                ps = [self._mk_bn(5 * r + i, getattr(rdb, f"conv{i}"), N, H, W, 4, Y, 24, 8 + 4 * i, out=D, out_planes=24,
                                  out_off=8 + 4 * i) for i in range(4)]
                dst, dplanes = (dense[r + 1], 24) if r + 1 < nr else (tr, 8)
                a, R = float(rdb.add_rate), float(rrdb.add_rate)
                if r % 3 < 2:      # add_rate * Conv(...) + rdb_in  (utils/models.py:271)
                    kw = dict(out_scale=a, res=D, res_planes=24, res_off=0, res_scale=1.0)
                else:              # R * (add_rate * Conv(...) + rdb_in) + rrdb_in  (utils/models.py:317)
                    kw = dict(out_scale=R * a, res=D, res_planes=24, res_off=0, res_scale=R, res2=dense[r - 2], res2_planes=24,
                              res2_off=0, res2_scale=1.0)
                ps.append(self._mk_bn(5 * r + 4, rdb.conv, N, H, W, 8, Y, 24, 0, out=dst, out_planes=dplanes, out_off=0, **kw))
                self.bn_pass.append(ps)
            # conv1 + long skip: t = x0 + BN(conv1(trunk))  (utils/models.py:616)
            self.bn_c1 = self._mk_bn(5 * nr, n.conv1, N, H, W, 8, rawc1, 8, 0, out=t, out_planes=8, out_off=0, res=dense[0],
                                     res_planes=24, res_off=0, res_scale=1.0)
        self.bn_sums.zero_()
        cv(self.f_conv0, xin, 1, 0, dense[0], 24, 0, N, H, W, slope=_slope_of(n.conv0.act), flags=ops.F_LRELU)
        for r, (rrdb, rdb, fwd, _, _) in enumerate(self.rdbs):
            D, Y, ps = dense[r], raw[r], self.bn_pass[r]
            for i in range(4):
                cv(fwd[i], D, 24, 0, Y, 24, 8 + 4 * i, N, H, W)
                ps[i].stats(self.device); ps[i].apply(self.device)
            cv(fwd[4], D, 24, 0, Y, 24, 0, N, H, W)
            ps[4].stats(self.device); ps[4].apply(self.device)
        cv(self.f_conv1, tr, 8, 0, rawc1, 8, 0, N, H, W)
        self.bn_c1.stats(self.device); self.bn_c1.apply(self.device)
        cur, h, w, ups = t, H, W, []
        for k, (s, fq) in enumerate(zip(n.scaler, self.f_scaler)):
            up = self._b(f"up{k}", N, 8, 2 * h, 2 * w)
            sl = _slope_of(s.net[2])
            for q in range(4):
                cv(fq[q], cur, 8, 0, up, 8, 2 * q, N, h, w, slope=sl, flags=ops.F_LRELU | ops.F_PIXSHUF2)
            ups.append((cur, up, h, w))
            cur, h, w = up, 2 * h, 2 * w
        out = torch.empty((N, 3, h, w), dtype=torch.float32, device=self.device)
        tanh = ops.F_TANH if isinstance(n.conv2.act, nn.Tanh) else 0
        cv(self.f_conv2, cur, 8, 0, out, 0, 0, N, h, w, flags=tanh | ops.F_F32OUT)
        self.saved = dict(N=N, H=H, W=W, xin=xin, dense=dense, raw=raw, rawc1=rawc1, tr=tr, t=t, ups=ups, out=out, tanh=bool(tanh))
        return out

    def _backward_impl(self, gout: torch.Tensor, run_batch: bool = True) -> list[torch.Tensor]:
        sv, n = self.saved, self.net
        if sv is None:
            raise RuntimeError("backward without a saved forward")
        N, H, W = sv["N"], sv["H"], sv["W"]
        cv = ops.conv2d_ex
        gv = self.grad_views
        self.flat_grad.zero_()
        self.bn_bsums.zero_()
        gkey = (N, H, W, tuple(sv["out"].shape))
        first = self.grad_batch is None or self.grad_key != gkey
        self._gjobs = [] if first else None
        gout = gout.float().contiguous()
        h2, w2 = sv["out"].shape[2], sv["out"].shape[3]
        if not sv["tanh"]:
            raise NotImplementedError("tail conv without tanh")
        sh, ish = float(self.grad_shift), 1.0 / float(self.grad_shift)
        # ---- tail conv and scalers: ConvWithoutBN, as TrainPlan
        g2 = ops.tanh_bwd_to_c8(gout, sv["out"], self.dt, scale=sh, g8=self._b("g2", N, 1, h2, w2))
        up_last = sv["ups"][-1][1] if sv["ups"] else sv["t"]
        ops.conv9_wgrad(up_last, 8, 0, 64, g2, 1, 0, 3, gv[id(n.conv2.conv.weight)], N, h2, w2, scale=ish)
        if self._gjobs is not None:
            self._gjobs.append(dict(dy=g2, dy_planes=1, dy_off=0, cout=3, db=gv[id(n.conv2.conv.bias)], N=N, H=h2, W=w2, scale=ish))
        g_cur = self._b("g_up_last", N, 8, h2, w2)
        cv(self.d_conv2, g2, 1, 0, g_cur, 8, 0, N, h2, w2)
        for k in range(len(sv["ups"]) - 1, -1, -1):
            cur_in, up, h, w = sv["ups"][k]
            s = n.scaler[k]
            g_s = self._b(f"g_s{k}", N, 32, h, w)
            ops.pixunshuffle_mask(g_cur, up, g_s, _slope_of(s.net[2]))
            self._wb(s.net[0], cur_in, 8, 0, 64, g_s, 32, 0, 256, N, h, w)
            g_in = self._b(f"g_in{k}", N, 8, h, w)
            cv(self.d_scaler[k], g_s, 32, 0, g_in, 8, 0, N, h, w)
            g_cur = g_in
        g_t = g_cur                                         # gradient of t = x0 + BN(conv1(trunk))
        nr = len(self.rdbs)
        Gd = [self._b(f"G{j}", N, 24, H, W) for j in range(nr)]     # [dY4 | dY3 | dY2 | dY1 | dY0]: gradients of the convs' RAW outputs
        GY = [self._b(f"GY{j}", N, 8, H, W) for j in range(nr)]     # gradient of the RDB's output (the residual term of the block in front)
        T = self._b("g_da", N, 4, H, W)                             # dL/da_k before the BatchNorm backward
        dyc1 = self._b("g_yc1", N, 8, H, W)
        gxf = self._b("g_x0t", N, 8, H, W)
        if first:
            self.bn_c1.set_backward(g_t, 8, 0, dyc1, 8, 0, gv[id(n.conv1.bn.weight)], gv[id(n.conv1.bn.bias)], 1.0, ish)
            for r, (rrdb, rdb, _, _, _) in enumerate(self.rdbs):
                a, R = float(rdb.add_rate), float(rrdb.add_rate)
                s_out = a * R if r % 3 == 2 else a
                ps, G = self.bn_pass[r], Gd[r]
                # the block's output scale stays folded into the transposed filters (TrainPlan._pack_jobs), so G[0..7] holds the
                # BatchNorm backward of the UNSCALED gy and only the parameter gradients carry s_out
                ps[4].set_backward(GY[r], 8, 0, G, 24, 0, gv[id(rdb.conv.bn.weight)], gv[id(rdb.conv.bn.bias)], 1.0, s_out * ish)
                for k in range(4):
                    m = getattr(rdb, f"conv{k}")
                    ps[k].set_backward(T, 4, 0, G, 24, 8 + 4 * (3 - k), gv[id(m.bn.weight)], gv[id(m.bn.bias)], 1.0, ish)
        # ---- conv1: BatchNorm backward, weight gradient, data gradient -> gy of the last RDB
        self.bn_c1.bwd_reduce(self.device); self.bn_c1.bwd_apply(self.device)
        self._wb(n.conv1, sv["tr"], 8, 0, 64, dyc1, 8, 0, 64, N, H, W)
        cv(self.d_conv1, dyc1, 8, 0, GY[nr - 1], 8, 0, N, H, W)
        for r in range(nr - 1, -1, -1):
            rrdb, rdb, _, stage, fin = self.rdbs[r]
            D, G, ps = sv["dense"][r], Gd[r], self.bn_pass[r]
            d = r % 3
            a, R = float(rdb.add_rate), float(rrdb.add_rate)
            s_out, s_res = (a * R, R) if d == 2 else (a, 1.0)
            ps[4].bwd_reduce(self.device); ps[4].bwd_apply(self.device)                  # GY[r] -> G[0..7]
            self._wb(rdb.conv, D, 24, 0, 192, G, 24, 0, 64, N, H, W, scale=s_out)
            for k in range(3, -1, -1):
                nin = 8 + 4 * (3 - k)
                cv(stage[k], G, 24, 0, T, 4, 0, N, H, W)                                 # dL/da_k from everything to its left
                ps[k].bwd_reduce(self.device); ps[k].bwd_apply(self.device)              # -> dY_k = G[nin .. nin+4)
                self._wb(getattr(rdb, f"conv{k}"), D, 24, 0, 64 + 32 * k, G, 24, nin, 32, N, H, W)
            dst = GY[r - 1] if r > 0 else gxf
            kw = dict(residual=GY[r], res_planes=8, res_off=0, res_scale=s_res)
            if d == 0:                                                                    # + the RRDB skip
                kw.update(res2=GY[r + 2], res2_planes=8, res2_off=0, res2_scale=1.0)
            cv(fin, G, 24, 0, dst, 8, 0, N, H, W, **kw)
        # ---- head conv
        gx0 = self._b("gx0", N, 8, H, W)
        ops.c8_axpby_mask(gxf, 8, 0, 1.0, g_t, 8, 0, 1.0, sv["dense"][0], 24, 0, _slope_of(n.conv0.act), gx0, 8, 0, 8, N, H, W)
        ops.conv9_wgrad(sv["xin"], 1, 0, 3, gx0, 8, 0, 64, gv[id(n.conv0.conv.weight)], N, H, W, scale=ish)
        if self._gjobs is not None:
            self._gjobs.append(dict(dy=gx0, dy_planes=8, dy_off=0, cout=64, db=gv[id(n.conv0.conv.bias)], N=N, H=H, W=W, scale=ish))
            self.grad_batch, self.grad_key = ops.GradBatch(self.ctx, self._gjobs, self.dt), gkey
            self.grad_parts = None
            self._gjobs = None
        self.grad_batch.run(self.device, self.wgrad_splitk)
        return [gv[id(p)] for p in self.params]

class _NetFunction(torch.autograd.Function):
    """preds = net(lr) with the whole forward / backward on the sm_100a kernels (train.py:55,57)."""

    @staticmethod
    def forward(ctx, x, plan, *params):
        ctx.plan = plan
        out = plan.forward(x)
        return out.view_as(out)     # fresh tensor object per step over the (graph-static) output buffer

    @staticmethod
    def backward(ctx, gout):
        grads = ctx.plan.backward(gout)
        # fresh tensor objects over the arena: autograd then adopts them as .grad (a tensor it does not own alone is cloned —
        # 488 device copies per step), and .grad stays a view of the flat arena for the optimizer
        return (None, None, *[g.detach() for g in grads])


def train_forward(net, x: torch.Tensor) -> torch.Tensor:
    plans = net.__dict__.setdefault("_train_plans", {})
    plan = plans.get(x.device)
    if plan is None:           # (also after deepcopy(model): see TrainPlan.__deepcopy__)
        bns = [m for m in net.modules() if isinstance(m, nn.BatchNorm2d)]
        cls = BNTrainPlan if bns else TrainPlan
        plan = plans[x.device] = cls(net, x.device, getattr(net, "train_operand_dtype", "fp16"))
    if plan.allow_bn and not all(m.training for m in plan.bn_modules):
        raise NotImplementedError("gradients through eval-mode BatchNorm (running statistics) are not on the sm_100a path: "
                                  "net.train() for training, torch.no_grad() for inference")
    return _NetFunction.apply(x.float().contiguous(), plan, *plan.params)


class _L1Function(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, tgt, w, squared=False):
        ctx.save_for_backward(pred, tgt)
        ctx.w, ctx.squared = w, squared
        return ops.l1_loss_fwd(pred, tgt, w, squared)

    @staticmethod
    def backward(ctx, g):
        pred, tgt = ctx.saved_tensors
        return ops.l1_loss_bwd_dev(pred, tgt, g, ctx.w, ctx.squared), None, None, None


class L1Loss(nn.Module):
    """Drop-in for utils/loss.py:32-39 (`sum(loss_weight * mean|x - y|)`) on the library's reduction kernels.
    The reference's default `L1Loss()` builds an int64 Parameter and raises on current torch; here the
    default is accepted and stored as float.  `loss_weight` is never given to an optimizer (train.py:265)."""

    def __init__(self, lossweight=1):
        super().__init__()
        self.loss_weight = nn.Parameter(torch.tensor([float(lossweight)]), requires_grad=False)
        self._w = (None, None)

    def _weight(self) -> float:
        v = self.loss_weight._version
        if self._w[0] != v:
            self._w = (v, float(self.loss_weight.detach().sum().item()))
        return self._w[1]

    def forward(self, inputs: torch.Tensor, ground_truth: torch.Tensor) -> torch.Tensor:
        if not inputs.is_cuda:
            raise RuntimeError("the B200 path has no CPU fallback: move predictions and targets to a CUDA device")
        pred = inputs.contiguous()
        tgt = ground_truth.float().contiguous()
        return _L1Function.apply(pred, tgt, self._weight())


class MSELoss(nn.Module):
    """torch.nn.MSELoss() as train.py uses it (train.py:13, :233, :264: the loss of a run without --enchant; default mean reduction)
    on the library's deterministic reduction kernels: loss = mean((pred - target)^2), d/dpred = 2 (pred - target) / numel."""

    def __init__(self, size_average=None, reduce=None, reduction: str = "mean"):
        super().__init__()
        if size_average is not None or reduce is not None or reduction != "mean":
            raise NotImplementedError("MSELoss: only the default mean reduction is on the sm_100a path (train.py:233)")
        self.reduction = reduction

    def forward(self, inputs: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        if not inputs.is_cuda:
            raise RuntimeError("the B200 path has no CPU fallback: move predictions and targets to a CUDA device")
        if inputs.shape != target.shape:
            raise RuntimeError(f"MSELoss: shapes differ: {tuple(inputs.shape)} vs {tuple(target.shape)}")
        return _L1Function.apply(inputs.contiguous(), target.float().contiguous(), 1.0, True)
