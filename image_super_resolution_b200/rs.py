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


def _window_groups(H: int, W: int, win: int):
    """The windows of rs.sliding_window(image, win) grouped by shape: [(y0, x0, ny, nx, h, w)] — ny x nx windows of h x w
    pixels starting at (y0, x0) with pitch (h, w).  Full windows, the right column, the bottom row and the corner."""
    fy, fx = H // win, W // win
    ry, rx = H - fy * win, W - fx * win
    groups = []
    if fy and fx:
        groups.append((0, 0, fy, fx, win, win))
    if fy and rx:
        groups.append((0, fx * win, fy, 1, win, rx))
    if ry and fx:
        groups.append((fy * win, 0, 1, fx, ry, win))
    if ry and rx:
        groups.append((fy * win, fx * win, 1, 1, ry, rx))
    return groups


@torch.no_grad()
def upscale_image(model: ModelPT | ShardedModel, image: torch.Tensor, window_size: int = 96, tile_mode: str = "reference",
                  tile: tuple[int, int] | None = None, batch: int = 256, out: torch.Tensor | None = None) -> torch.Tensor:
    """uint8 [3,H,W] (any device) -> uint8 [3,H,W] on the CPU, like rs.py:82-111.

    The image is uploaded once per device and the result assembled on the device (one download at the end); the windows /
    tiles are device-side slices.  tile_mode="reference": the non-overlapping windows of rs.py:90-111, each forwarded as its
    own image — windows of equal shape ride in one batch (`batch` per forward: a window's result does not depend on its
    batch mates), so the seams are the reference's while the launches are not one per 96 x 96 window.
    out: optional uint8 [3,H,W] CPU tensor that receives the result (a reused — ideally pinned — buffer spares the 20 ms of
    page faults a fresh 48 MB host tensor costs for a 4096 x 4096 image)."""
    if image.dim() != 3 or image.dtype != torch.uint8:
        raise RuntimeError("expected a uint8 [3,H,W] image (torchvision.io.read_image output)")
    sm = model if isinstance(model, ShardedModel) else None
    reps = sm.replicas if sm else [model]
    G = len(reps)
    _, H, W = image.shape
    if tile_mode not in ("whole", "reference", "halo"):
        raise ValueError(f"tile_mode {tile_mode!r}")
    if out is not None and (out.is_cuda or out.dtype != torch.uint8 or tuple(out.shape) != (3, H, W) or not out.is_contiguous()):
        raise RuntimeError("out must be a contiguous uint8 [3,H,W] CPU tensor")

    def download(t: torch.Tensor) -> torch.Tensor:
        if out is None:
            r = t.cpu()
        else:
            out.copy_(t, non_blocking=True)
            torch.cuda.current_stream(t.device).synchronize()
            r = out
        for m in reps:
            m._ctx.poll()          # a kernel-side pipeline timeout raises here instead of returning garbage pixels
        return r

    src = image if not image.is_cuda else image.cpu()
    if G == 1 and image.is_cuda and image.device == reps[0]._device:
        imgs = [image.contiguous()]
    else:
        imgs = [src.to(m._device, non_blocking=True) for m in reps]          # one upload per device
    canvases = [torch.zeros_like(t) for t in imgs]                            # every unit pastes into its device's canvas
    rects = [[] for _ in range(G)]                                            # what each device's canvas holds
    unit = 0

    def run(x, paste):
        nonlocal unit
        d = unit % G
        unit += 1
        with torch.cuda.device(reps[d]._device):
            paste(d, reps[d].forward(x(imgs[d])))

    if tile_mode == "whole":
        with torch.cuda.device(reps[0]._device):
            return download(reps[0].forward(imgs[0].unsqueeze(0))[0])
    if tile_mode == "reference":
        for y0, x0, ny, nx, h, w in _window_groups(H, W, window_size):
            rows_per = max(1, batch // nx)                                    # whole rows of windows per forward
            for r0 in range(0, ny, rows_per):
                r1 = min(ny, r0 + rows_per)
                ya, yb = y0 + r0 * h, y0 + r1 * h

                def gather(img, ya=ya, yb=yb, x0=x0, nx=nx, h=h, w=w, nr=r1 - r0):
                    blk = img[:, ya:yb, x0:x0 + nx * w].reshape(3, nr, h, nx, w)
                    return blk.permute(1, 3, 0, 2, 4).reshape(nr * nx, 3, h, w).contiguous()

                def paste(d, out, ya=ya, yb=yb, x0=x0, nx=nx, h=h, w=w, nr=r1 - r0):
                    canvases[d][:, ya:yb, x0:x0 + nx * w] = out.reshape(nr, nx, 3, h, w).permute(2, 0, 3, 1, 4).reshape(3, nr * h, nx * w)
                    rects[d].append((ya, yb, x0, x0 + nx * w))

                run(gather, paste)
    else:
        th, tw = tile if tile is not None else (max(window_size, 1024), max(window_size, 1024))
        for y, x, h, w, y0, x0, y1, x1 in plan_tiles(H, W, th, tw, RF_RADIUS):
            def crop(img, y0=y0, y1=y1, x0=x0, x1=x1):
                return img[:, y0:y1, x0:x1].unsqueeze(0).contiguous()

            def paste(d, out, y=y, x=x, h=h, w=w, y0=y0, x0=x0):
                canvases[d][:, y:y + h, x:x + w] = out[0, :, y - y0:y - y0 + h, x - x0:x - x0 + w]
                rects[d].append((y, y + h, x, x + w))

            run(crop, paste)
    if G == 1:
        with torch.cuda.device(reps[0]._device):
            return download(canvases[0])
    result = out if out is not None else torch.zeros((3, H, W), dtype=torch.uint8)
    for d, c in enumerate(canvases):                                           # merge the devices' canvases on the host
        if rects[d]:
            ch = c.cpu()
            for ya, yb, xa, xb in rects[d]:
                result[:, ya:yb, xa:xb] = ch[:, ya:yb, xa:xb]
    return result


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
