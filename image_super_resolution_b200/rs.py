"""Drop-in for the reference's inference entry `rs.py` (sliding_window :16-27, runer :39-114).

Kept: `sliding_window(image, step, windowSize=None)` (same generator), `runer(**kwargs)` with the keys
`model, src, save_dir, window_size, batch_size, worker`.  Added (additive keyword arguments):
  tile_mode = "reference"  non-overlapping windows, each forwarded on its own — seam-for-seam what
                           rs.py:90-111 produces;
            = "halo"       every tile extended by the receptive-field radius (25 px) and clipped at the image
                           border: bit-identical to the whole-image forward;
            = "whole"      one forward over the whole image;
  devices   = list of CUDA device indices; tiles / frames are dealt round-robin, no inter-GPU traffic.
File decode/encode stays the caller's (torchvision) business; the video writer is a headless byte sink with
`FFMPEG_recorder.writeFrame`'s signature (utils/ffmpeg.py:69-76) because no `ffmpeg` binary exists here.
"""
from __future__ import annotations

import collections
from pathlib import Path
from typing import Iterable, Iterator, Sequence

import torch

from .modelpt import ModelPT, RF_RADIUS, load


def sliding_window(image: torch.Tensor, step, windowSize=None):
    """rs.py:16-27 — accepts ...CHW; yields (step, x, y, view) over non-overlapping windows."""
    if windowSize is None:
        windowSize = step
    if isinstance(step, int):
        step = [step] * 2
    step[0] = min(image.shape[-2], step[0])
    step[1] = min(image.shape[-1], step[1])
    for y in range(0, image.shape[-2], step[0]):
        for x in range(0, image.shape[-1], step[1]):
            yield step, x, y, image[..., y:y + windowSize, x:x + windowSize]


def total_inter_sliding_window(image: torch.Tensor, step: int) -> int:
    """rs.py:30-36."""
    if image.dim() == 3:
        _, h, w = image.size()
        return len(range(0, h, step)) * len(range(0, w, step))
    return 0


# ------------------------------------------------------------------------------------------------
# tile plans (pure host logic, unit-tested on CPU)
# ------------------------------------------------------------------------------------------------
def plan_tiles(H: int, W: int, tile_h: int, tile_w: int, halo: int):
    """[(y, x, th, tw, y0, x0, y1, x1)]: interior [y:y+th, x:x+tw] computed from the clipped window [y0:y1, x0:x1]."""
    plan = []
    for y in range(0, H, tile_h):
        for x in range(0, W, tile_w):
            th, tw = min(tile_h, H - y), min(tile_w, W - x)
            y0, x0 = max(0, y - halo), max(0, x - halo)
            y1, x1 = min(H, y + th + halo), min(W, x + tw + halo)
            plan.append((y, x, th, tw, y0, x0, y1, x1))
    return plan


def shard(n_items: int, rank: int, world: int) -> range:
    """Item i goes to rank i mod world (frames / tiles are independent units)."""
    return range(rank, n_items, world)


class ShardedModel:
    """One ModelPT replica per device; call(i, x) runs unit i on device i mod G, asynchronously."""

    def __init__(self, model: ModelPT | str, devices: Sequence[int] | None = None):
        if devices is None:
            devices = [torch.cuda.current_device()]
        self.devices = list(devices)
        if isinstance(model, (str, Path)):
            state = load(model).state_dict()
        else:
            state = model.state_dict()
        self.replicas = [ModelPT(state).to(torch.device("cuda", d)).eval() for d in self.devices]

    def __len__(self):
        return len(self.devices)

    def forward_on(self, slot: int, x: torch.Tensor, layout: str = "nchw") -> torch.Tensor:
        d = self.devices[slot % len(self.devices)]
        m = self.replicas[slot % len(self.devices)]
        with torch.cuda.device(d):
            return m.forward(x.to(torch.device("cuda", d), non_blocking=True), layout=layout)


@torch.no_grad()
def upscale_image(model: ModelPT | ShardedModel, image: torch.Tensor, window_size: int = 96, tile_mode: str = "reference",
                  tile: tuple[int, int] | None = None) -> torch.Tensor:
    """uint8 [3,H,W] (any device) -> uint8 [3,H,W] on the CPU, like rs.py:82-111."""
    if image.dim() != 3 or image.dtype != torch.uint8:
        raise RuntimeError("expected a uint8 [3,H,W] image (torchvision.io.read_image output)")
    sm = model if isinstance(model, ShardedModel) else None
    run = (lambda i, x: sm.forward_on(i, x)) if sm else (lambda i, x: model(x.to(model._device, non_blocking=True)))
    _, H, W = image.shape
    result = torch.zeros_like(image, device="cpu")
    if tile_mode == "whole":
        return run(0, image.unsqueeze(0))[0].cpu()
    if tile_mode == "reference":
        pending = []
        for i, (_, x, y, win) in enumerate(sliding_window(image, window_size)):
            pending.append((y, x, run(i, win.unsqueeze(0).contiguous())))
        for y, x, out in pending:  # paste in raster order, like rs.py:105-111
            o = out[0].cpu()
            result[:, y:y + o.shape[-2], x:x + o.shape[-1]] = o
        return result
    if tile_mode == "halo":
        th, tw = tile if tile is not None else (max(window_size, 256), max(window_size, 256))
        pending = []
        for i, (y, x, h, w, y0, x0, y1, x1) in enumerate(plan_tiles(H, W, th, tw, RF_RADIUS)):
            pending.append((y, x, h, w, y0, x0, run(i, image[:, y0:y1, x0:x1].unsqueeze(0).contiguous())))
        for y, x, h, w, y0, x0, out in pending:
            result[:, y:y + h, x:x + w] = out[0, :, y - y0:y - y0 + h, x - x0:x - x0 + w].cpu()
        return result
    raise ValueError(f"tile_mode {tile_mode!r}")


class ByteSink:
    """Headless stand-in for FFMPEG_recorder (utils/ffmpeg.py:28-140): same writeFrame(ndarray HWC uint8 BGR)."""

    def __init__(self, keep: bool = False):
        self.frames, self.bytes, self.keep, self.kept = 0, 0, keep, []

    def writeFrame(self, frame):
        b = frame.tobytes()
        self.frames += 1
        self.bytes += len(b)
        if self.keep:
            self.kept.append(frame.copy())

    def stopRecorder(self):
        return None

    def addAudio(self, *_a, **_k):
        return None


@torch.no_grad()
def upscale_frames(model: ModelPT | ShardedModel, batches: Iterable[torch.Tensor], sink, bgr: bool = True) -> int:
    """rs.py:60-74 with the model call replaced: each uint8 [B,3,H,W] batch is one unit, dealt round-robin over the
    devices; results leave in frame order as HWC (BGR like RGB2BGR, utils/datasets.py:113-124) into sink.writeFrame."""
    sm = model if isinstance(model, ShardedModel) else None
    G = len(sm) if sm else 1
    layout = "hwc_bgr" if bgr else "nchw"               # BGR frames: the tail conv stores HWC / BGR itself (no glue kernels)
    n = 0

    def emit(o):
        nonlocal n
        for f in o:
            sink.writeFrame(f)
            n += 1

    # Every device runs the pipelined host-to-host stream of its replica (H2D of the next batch and D2H of the previous one
    # overlap the kernels; the forward is a captured graph); batch i goes to device i mod G and the results are taken in
    # frame order, so the writer gets each pinned result while the following batches compute.
    reps = sm.replicas if sm else [model]
    src = iter(batches)
    queues = [collections.deque() for _ in range(G)]
    state = {"next": 0, "done": False}

    def feeder(d):
        """Batches d, d + G, ... of the shared source (pulled lazily, in order)."""
        while True:
            while not queues[d] and not state["done"]:
                try:
                    b = next(src)
                except StopIteration:
                    state["done"] = True
                    break
                queues[state["next"] % G].append(b if not b.is_cuda else b.cpu())
                state["next"] += 1
            if not queues[d]:
                return
            yield queues[d].popleft()

    gens = [m.stream(feeder(d), depth=2, layout=layout) for d, m in enumerate(reps)]
    i = 0
    while True:
        m = reps[i % G]
        with torch.cuda.device(m._device):
            y = next(gens[i % G], None)
        if y is None:
            break
        emit((y if bgr else y.permute(0, 2, 3, 1).contiguous()).numpy())
        i += 1
    return n


def runer(**kwargs):
    """rs.py:39-114 with the same keyword arguments (plus tile_mode / devices)."""
    from torchvision.io import read_image, ImageReadMode, write_png
    model_dir, src, result_dir = Path(kwargs["model"]), Path(kwargs["src"]), Path(kwargs["save_dir"])
    step_size = kwargs["window_size"]
    tile_mode = kwargs.get("tile_mode", "reference")
    devices = kwargs.get("devices")
    if not torch.cuda.is_available():
        raise RuntimeError("runer: no CUDA device — this build has no CPU path")
    model = ShardedModel(str(model_dir), devices) if devices else load(str(model_dir)).to("cuda").eval()
    image = read_image(src.as_posix(), ImageReadMode.RGB)
    print("input shape", image.size())
    result = upscale_image(model, image, step_size, tile_mode)
    result_dir = result_dir.with_suffix(".png")
    write_png(result, result_dir.as_posix())
    print("output shape", result.shape, f"{result_dir.as_posix()}")
    return result
