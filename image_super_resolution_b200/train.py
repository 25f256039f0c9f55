"""Drop-in for the reference's pixel-loss training driver (train.py:41-67 `train`, utils/models.py:17-40 `ModelEMA`).

`train(model, ema, dataloader, compute_loss, optimizer, gradscaler, schedule, epoch, tensorBoard)` keeps the reference's
signature, call order (zero_grad -> autocast forward -> loss -> scaled backward -> unscale_ -> clip_grad_norm_(10) ->
step -> update -> schedule.step -> ema.update -> log) and return value (list of per-step losses).  With the drop-in
`EResNet` and `L1Loss` the forward, loss and backward run on the sm_100a kernels (training.py); optimizer, GradScaler
and LR schedule stay the caller's PyTorch objects, exactly as in the reference.

Data parallelism (new; the reference is single-process): one process per GPU, `init_distributed()` joins the
NCCL group from the torchrun environment, every rank feeds its own shard of the global batch and the flat gradient
arena is all-reduced (mean) inside backward — between train.py:57 and :58 — by training.TrainPlan.  Rank 0 logs.
"""
from __future__ import annotations

import math
import os
from copy import deepcopy

import torch
import torch.distributed as dist

from .optim import clip_grad_norm_, FlatAdam, GradScaler  # noqa: F401  (torch's own for anything but a FlatAdam)


class ModelEMA:
    """utils/models.py:17-40: EMA over the whole state_dict, decay ramp d = decay * (1 - exp(-updates / tau)).
    Same arithmetic per tensor (v = d*v + (1-d)*m); issued as two multi-tensor launches instead of a per-tensor loop."""

    def __init__(self, model, decay=0.9999, tau=2000, updates=0):
        self.ema = deepcopy(model).eval()
        self.updates = updates
        self.decay = lambda x: decay * (1 - math.exp(-x / tau))
        for p in self.ema.parameters():
            p.requires_grad = False
        self._opt = None

    def attach(self, optimizer):
        """Let `optimizer` (optim.FlatAdam over this model's parameters) blend the EMA parameters in its own pass over the
        parameter arena; `update()` then only advances the counter (and blends floating-point buffers, if any)."""
        ps = [p for p in self.ema.parameters()]
        if any(not p.is_cuda for p in ps):
            raise RuntimeError("ModelEMA.attach: move the EMA model to the optimizer's device first (ema.ema.to(device))")
        from .optim import flatten_
        with torch.no_grad():
            arena = flatten_(ps)
        optimizer.attach_ema(self, arena)
        self._opt = optimizer
        return self

    @torch.no_grad()
    def update(self, model):
        self.updates += 1
        d = self.decay(self.updates)
        # the floating-point entries of the two state_dicts, paired in module order (parameters, then buffers) — the same
        # tensors `state_dict()` would return, without building 2 x 488 prefixed keys every step (4 ms of host time)
        dst, src = [], []
        if self._opt is not None and self._opt.ema_applied_for == self.updates:      # parameters: done by the optimizer's pass
            # (the buffer pairs are cached per model: walking two module trees of 1 000 modules costs the host 1 ms a step)
            c = self.__dict__.get("_buf_pairs")
            if c is None or c[0] is not model:
                c = self.__dict__["_buf_pairs"] = (model, list(zip(list(self.ema.buffers()), list(model.buffers()))))
            pairs = c[1]
        else:
            pairs = zip(list(self.ema.parameters()) + list(self.ema.buffers()), list(model.parameters()) + list(model.buffers()))
        for v, m in pairs:
            if v.dtype.is_floating_point:
                dst.append(v)
                src.append(m.detach())
        if dst:
            torch._foreach_mul_(dst, d)
            torch._foreach_add_(dst, src, alpha=1 - d)


def init_distributed(backend: str | None = None) -> tuple[int, int, torch.device]:
    """Join the process group described by torchrun's environment (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_*).
    Returns (rank, world_size, device).  Single-process runs return (0, 1, cuda:0 or cpu) without a group."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    cuda = torch.cuda.is_available()
    device = torch.device("cuda", local) if cuda else torch.device("cpu")
    if cuda:
        torch.cuda.set_device(device)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        dist.init_process_group(backend or ("nccl" if cuda else "gloo"), rank=rank, world_size=world,
                                **({"device_id": device} if cuda else {}))
    return rank, world, device


def shard_of(batch: torch.Tensor, rank: int, world: int) -> torch.Tensor:
    """This rank's contiguous slice of a global batch (global batch = per-GPU batch x world)."""
    n = batch.shape[0]
    if n % world:
        raise ValueError(f"global batch {n} is not divisible by world size {world}")
    per = n // world
    return batch[rank * per:(rank + 1) * per]


def broadcast_parameters(model, src: int = 0):
    """Replicas start from rank `src`'s weights (what DDP's constructor does)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        for t in list(model.parameters()) + list(model.buffers()):
            dist.broadcast(t.data, src)
        from . import training
        training.PARAM_EPOCH[0] += 1            # written through .data: autograd's version counters did not move


class _NullBoard:
    def add_scalar(self, *a, **k):
        pass


def train(model, ema, dataloader, compute_loss, optimizer, gradscaler, schedule, epoch: int, tensorBoard=None):
    """train.py:41-67.  Returns the list of per-step loss values (python floats), like the reference."""
    model.train()
    losses = []
    device = next(model.parameters()).device
    if ema is not None:
        ema.ema.to(device)
    autocast_device = "cuda" if device.type == "cuda" else "cpu"
    rank0 = not (dist.is_available() and dist.is_initialized()) or dist.get_rank() == 0
    board = tensorBoard if (tensorBoard is not None and rank0) else _NullBoard()
    total = len(dataloader)
    deferred = [] if (isinstance(board, _NullBoard) and device.type == "cuda") else None
    for idx, (hr_images, lr_images) in enumerate(dataloader):
        hr_images, lr_images = hr_images.to(device, non_blocking=True), lr_images.to(device, non_blocking=True)
        optimizer.zero_grad()
        with torch.autocast(enabled=device.type == "cuda", device_type=autocast_device):
            preds = model(lr_images)
            loss = compute_loss(preds, hr_images)
        gradscaler.scale(loss).backward()          # gradients are all-reduced inside backward when world_size > 1
        gradscaler.unscale_(optimizer)
        clip_grad_norm_(model.parameters(), 10)
        gradscaler.step(optimizer)
        gradscaler.update()
        if schedule is not None:
            schedule.step()
        if ema is not None:
            ema.update(model)
        if deferred is None:
            value = loss.item()                    # the reference reads it back every step (train.py:64-65) for its logger
            losses.append(value)
            board.add_scalar("loss", value, epoch * total + idx + 1)
        else:
            deferred.append(loss.detach().clone())
    if deferred:
        # nobody watches the per-step value (no tensorBoard): the read-backs are taken together at the end of the epoch, so
        # the host keeps enqueueing the next step while the device finishes this one (a .item() per step leaves the GPU idle
        # for the host's whole launch path: 0.7 of 9.7 ms).  Same return value: the list of per-step losses.
        losses = torch.stack(deferred).float().cpu().tolist()
    if device.type == "cuda":
        from ._lib import Context
        Context.get(device.index or 0).poll()      # a kernel-side pipeline timeout during the epoch raises here
    return losses
