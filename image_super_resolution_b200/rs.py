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


class _Lane:
    """Per-replica tile pipeline: upload / compute / download streams and two slots of device buffers, so the transfer of a
    unit overlaps the kernels of its neighbours on the same device."""

    def __init__(self, m: ModelPT):
        self.m, self.dev = m, m._device
        self.s_in, self.s_c, self.s_out = (torch.cuda.Stream(self.dev) for _ in range(3))
        self.slots = [dict(x=None, y=None, ev_in=torch.cuda.Event(), ev_c=torch.cuda.Event(), ev_out=torch.cuda.Event())
                      for _ in range(2)]
        self.k = 0

    @staticmethod
    def _buf(slot, key, nbytes, dev):
        t = slot[key]
        if t is None or t.numel() < nbytes:
            slot[key] = t = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        return t


def _lanes(model) -> list:
    reps = model.replicas if isinstance(model, ShardedModel) else [model]
    lanes = []
    for m in reps:
        ln = m.__dict__.get("_rs_lane")
        if ln is None or ln.dev != m._device:
            ln = m.__dict__["_rs_lane"] = _Lane(m)
        lanes.append(ln)
    return lanes


def _rect_copy(ctx, dst_ptr, dst_pitch, dst_plane, src_ptr, src_pitch, src_plane, w, h, kind, stream):
    import ctypes as C
    from ._lib import lib, check
    check(lib().srb_copy_rect_u8(ctx.handle, C.c_void_p(dst_ptr), dst_pitch, dst_plane, C.c_void_p(src_ptr), src_pitch, src_plane, 3,
                                 w, h, kind, C.c_void_p(stream.cuda_stream)), "srb_copy_rect_u8")


@torch.no_grad()
def upscale_image(model: ModelPT | ShardedModel, image: torch.Tensor, window_size: int = 96, tile_mode: str = "reference",
                  tile: tuple[int, int] | None = None, batch: int = 256, out: torch.Tensor | None = None) -> torch.Tensor:
    """uint8 [3,H,W] -> uint8 [3,H,W] on the CPU, like rs.py:82-111.

    The image is cut into UNITS (tile_mode="halo": a tile with its receptive-field halo; "reference": a band of rs.py's
    non-overlapping windows, windows of equal shape riding in one batch — a window's result does not depend on its batch
    mates, so the seams are the reference's; "whole": the image).  Unit u runs on device u mod G.  Only the unit's rectangle
    is uploaded (2-D copies straight out of the pinned image) and only its interior is downloaded, straight into the stitched
    result — no whole-image transfers, no host-side crop / paste copies; on every device the transfers of a unit overlap the
    kernels of the next one (three streams, two slots).  No inter-GPU traffic.
    out: optional uint8 [3,H,W] CPU tensor that receives the result (pinned memory keeps the downloads asynchronous; a fresh
    pinned result is allocated otherwise).  A pinned `image` is read in place; a pageable one is first copied to pinned memory."""
    if image.dim() != 3 or image.dtype != torch.uint8 or image.shape[0] != 3:
        raise RuntimeError("expected a uint8 [3,H,W] image (torchvision.io.read_image output)")
    if tile_mode not in ("whole", "reference", "halo"):
        raise ValueError(f"tile_mode {tile_mode!r}")
    _, H, W = image.shape
    if out is not None and (out.is_cuda or out.dtype != torch.uint8 or tuple(out.shape) != (3, H, W) or not out.is_contiguous()):
        raise RuntimeError("out must be a contiguous uint8 [3,H,W] CPU tensor")
    lanes = _lanes(model)
    G = len(lanes)
    if image.is_cuda or not image.is_contiguous() or not image.is_pinned():
        src = torch.empty((3, H, W), dtype=torch.uint8, pin_memory=True)
        src.copy_(image)
    else:
        src = image
    result = out if out is not None else torch.empty((3, H, W), dtype=torch.uint8, pin_memory=True)
    sp, rp = src.data_ptr(), result.data_ptr()

    # ---- units: (input rectangle, fn(device rect [3,h,w], y buffer) -> device result [3,oh,ow], (ry, rx, oh', ow', dst y, dst x))
    units = []
    if tile_mode == "whole":
        units.append(((0, H, 0, W), None, (0, 0, H, W, 0, 0)))
    elif tile_mode == "halo":
        th, tw = tile if tile is not None else (max(window_size, 1024), max(window_size, 1024))
        for y, x, h, w, y0, x0, y1, x1 in plan_tiles(H, W, th, tw, RF_RADIUS):
            units.append(((y0, y1, x0, x1), None, (y - y0, x - x0, h, w, y, x)))
    else:
        for gy0, gx0, ny, nx, h, w in _window_groups(H, W, window_size):
            rows_per = max(1, batch // nx)                                    # whole rows of windows per forward
            for r0 in range(0, ny, rows_per):
                nr = min(ny, r0 + rows_per) - r0
                ya = gy0 + r0 * h
                units.append(((ya, ya + nr * h, gx0, gx0 + nx * w), (nr, nx, h, w), (0, 0, nr * h, nx * w, ya, gx0)))

    for ln in lanes[:min(G, len(units))]:
        ln.s_c.wait_stream(torch.cuda.current_stream(ln.dev))   # the replica's workspace may be in use by the caller's stream
    for u, ((y0, y1, x0, x1), windows, (ry, rx, oh, ow, dy, dx)) in enumerate(units):
        ln = lanes[u % G]
        m, ctx = ln.m, ln.m._ctx
        sl = ln.slots[ln.k % 2]
        ln.k += 1
        h, w = y1 - y0, x1 - x0
        with torch.cuda.device(ln.dev):
            xb = _Lane._buf(sl, "x", 3 * h * w, ln.dev)[:3 * h * w].view(3, h, w)
            yb = _Lane._buf(sl, "y", 3 * h * w, ln.dev)[:3 * h * w]
            ln.s_in.wait_event(sl["ev_c"])                       # the kernels that last read this slot's x are done
            _rect_copy(ctx, xb.data_ptr(), w, h * w, sp + y0 * W + x0, W, H * W, w, h, 0, ln.s_in)
            sl["ev_in"].record(ln.s_in)
            ln.s_c.wait_event(sl["ev_in"])
            ln.s_c.wait_event(sl["ev_out"])                      # this slot's previous result has left the device
            with torch.cuda.stream(ln.s_c):
                if windows is None:
                    y_dev = m.forward(xb.unsqueeze(0), out=yb.view(1, 3, h, w))[0]
                else:                                            # rs.py:90-111: every window forwarded as its own image
                    nr, nx, wh, ww = windows
                    xs = xb.reshape(3, nr, wh, nx, ww).permute(1, 3, 0, 2, 4).reshape(nr * nx, 3, wh, ww).contiguous()
                    o = m.forward(xs)
                    y_dev = yb.view(3, h, w)
                    y_dev.copy_(o.reshape(nr, nx, 3, wh, ww).permute(2, 0, 3, 1, 4).reshape(3, h, w))
                if m._ws is not None:
                    m._ws.record_stream(ln.s_c)                  # (the workspace may have been allocated under another stream)
                sl["ev_c"].record(ln.s_c)
            ln.s_out.wait_event(sl["ev_c"])
            _rect_copy(ctx, rp + dy * W + dx, W, H * W, y_dev.data_ptr() + ry * w + rx, w, h * w, ow, oh, 1, ln.s_out)
            sl["ev_out"].record(ln.s_out)
    for ln in lanes[:min(G, len(units))]:
        ln.s_out.synchronize()
        ln.m._ctx.poll()           # a kernel-side pipeline timeout raises here instead of returning garbage pixels
    return result


class ByteSink:
    """Headless stand-in for FFMPEG_recorder (utils/ffmpeg.py:28-140): same writeFrame(ndarray HWC uint8 BGR)."""

    def __init__(self, keep: bool = False, copy: bool = True):
        """copy=True: `frame.tobytes()` like the reference's writer (utils/ffmpeg.py:74: one host memcpy of the frame into the
        pipe of the encoder process); copy=False only counts the bytes — for measuring this library's path without the cost
        of the stand-in for the external encoder (6.2 MB memcpy per 1080p frame: one host thread tops out near 1 500 fps)."""
        self.frames, self.bytes, self.keep, self.kept, self.copy = 0, 0, keep, [], copy

    def writeFrame(self, frame):
        b = frame.tobytes() if self.copy else memoryview(frame).cast("B") if frame.flags["C_CONTIGUOUS"] else frame.tobytes()
        self.frames += 1
        self.bytes += len(b)
        if self.keep:
            self.kept.append(frame.copy())

    def stopRecorder(self):
        return None

    def addAudio(self, *_a, **_k):
        return None


@torch.no_grad()
def upscale_frames(model: ModelPT | ShardedModel, batches: Iterable[torch.Tensor], sink, bgr: bool = True,
                   threads: bool | None = None, in_layout: str = "nchw") -> int:
    """rs.py:60-74 with the model call replaced: each uint8 [B,3,H,W] batch is one unit, dealt round-robin over the
    devices; results leave in frame order as HWC (BGR like RGB2BGR, utils/datasets.py:113-124) into sink.writeFrame.

    Every device runs the pipelined host-to-host stream of its replica (H2D of the next batch and D2H of the previous one
    overlap the kernels; the forward is a captured graph); batch i goes to device i mod G and the results are taken in frame
    order, so the writer gets each pinned result while the following batches compute.  threads (default: G > 1): one feeder
    thread per device drives its replica's stream — the CUDA calls release the GIL, so the enqueue work of G devices and the
    writer's memcpy overlap instead of queueing behind one Python thread — and the calling thread only writes frames.
    in_layout="hwc_rgb" / "hwc_bgr": the batches are [B,H,W,3] frames as a decoder delivers them; they are uploaded as they
    are and conv0's loader reads them (no permute, `.float()` or Normalize pass: rs.py:60-63 disappears into the head kernel)."""
    import threading
    import queue as _queue
    sm = model if isinstance(model, ShardedModel) else None
    G = len(sm) if sm else 1
    layout = "hwc_bgr" if bgr else "nchw"               # BGR frames: the tail conv stores HWC / BGR itself (no glue kernels)
    if threads is None:
        threads = G > 1
    n = 0

    def emit(y):
        nonlocal n
        for f in (y if bgr else y.permute(0, 2, 3, 1).contiguous()).numpy():
            sink.writeFrame(f)
            n += 1

    reps = sm.replicas if sm else [model]
    src = iter(batches)
    queues = [collections.deque() for _ in range(G)]
    state = {"next": 0, "done": False}
    lock = threading.Lock()

    def feeder(d):
        """Batches d, d + G, ... of the shared source (pulled lazily, in order)."""
        while True:
            with lock:
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
                item = queues[d].popleft()
            yield item

    if not threads:
        gens = [m.stream(feeder(d), depth=2, layout=layout, in_layout=in_layout) for d, m in enumerate(reps)]
        i = 0
        while True:
            m = reps[i % G]
            with torch.cuda.device(m._device):
                y = next(gens[i % G], None)
            if y is None:
                break
            emit(y)
            i += 1
        return n

    # a yielded pinned tensor stays valid until `depth` more items have been taken from its generator (ModelPT.stream): with a
    # one-place hand-over queue the worker is at most two items ahead of the frame being written
    outq = [_queue.Queue(maxsize=1) for _ in range(G)]
    stop = threading.Event()

    def worker(d):
        try:
            with torch.cuda.device(reps[d]._device):
                for y in reps[d].stream(feeder(d), depth=2, layout=layout, in_layout=in_layout):
                    while not stop.is_set():
                        try:
                            outq[d].put(y, timeout=0.1)
                            break
                        except _queue.Full:
                            pass
                    if stop.is_set():
                        return
            outq[d].put(None)
        except BaseException as e:  # noqa: BLE001  (handed to the calling thread)
            outq[d].put(e)

    ths = [threading.Thread(target=worker, args=(d,), daemon=True) for d in range(G)]
    for t in ths:
        t.start()
    try:
        i = 0
        while True:
            y = outq[i % G].get()
            if y is None:
                break
            if isinstance(y, BaseException):
                raise y
            emit(y)
            i += 1
    finally:
        stop.set()
        for t in ths:
            t.join(timeout=5)
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
