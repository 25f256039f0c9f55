"""Drop-in for the object `torch.jit.load("model.pt")` returns in the reference (rs.py:48-49).

    model = image_super_resolution_b200.load("model.pt")     # or the extracted .npz
    model.to("cuda").eval()
    out_u8 = model(window_u8)          # rs.py:94  uint8 [N,3,H,W] -> uint8 [N,3,H,W]

The graph (Normalize -> conv0 -> 8 x ResidualBlock1 -> conv1 + skip -> conv2 + tanh -> uint8;
SURVEY §8(a) a1-a7) runs entirely inside libsrb200.so on sm_100a; this class only owns the
packed weights and a workspace tensor.  There is no PyTorch / CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from collections import OrderedDict

import numpy as np
import torch

from . import _lib
from ._lib import Context, check, lib
from .ops import dtype_code

# state-dict keys of the 19 convs in graph order (archive __torch__.py, class Denoise)
CONV_KEYS = (["net.1.conv0.0.conv"] + [f"net.1.residual.{i}.m.{j}.conv" for i in range(8) for j in (0, 1)] +
             ["net.1.conv1.conv", "net.1.conv2.0.conv"])
LRELU_SLOPE = 0.2          # leaky_relu_(0.2) in the traced graph
RF_RADIUS = 25             # receptive-field radius of the graph
FLOP_PER_PX = 725760       # algorithmic FLOP per pixel


def _load_state(path: str) -> "OrderedDict[str, torch.Tensor]":
    path = str(path)
    if path.endswith(".npz"):
        z = np.load(path)
        return OrderedDict((k, torch.from_numpy(z[k].copy())) for k in z.files)
    m = torch.jit.load(path, "cpu")  # exactly as rs.py:48 opens it (TorchScript archive)
    return OrderedDict((k, v.detach().clone()) for k, v in m.state_dict().items() if "store_bn" not in k)


class ModelPT:
    """Callable with the surface rs.py uses: .to(device), .eval(), __call__(uint8 NCHW), .state_dict()."""

    def __init__(self, state: "OrderedDict[str, torch.Tensor]", dtype="fp16"):
        missing = [k + s for k in CONV_KEYS for s in (".weight", ".bias") if k + s not in state]
        if missing:
            raise KeyError(f"not a model.pt state dict, missing {missing[:3]}...")
        self._state = state
        self._dtype = dtype_code(dtype)
        self._device = None
        self._net = None
        self._ctx = None
        self._ws = None
        self.training = False

    # ---- torch.nn.Module-like surface -------------------------------------------------------
    def state_dict(self):
        return OrderedDict(self._state)

    def eval(self):
        self.training = False
        return self

    def train(self, mode: bool = True):
        if mode:
            raise RuntimeError("model.pt is an inference artefact (TorchScript trace); training uses the class nets")
        return self

    def to(self, device):
        device = torch.device(device)
        if device.type != "cuda":
            raise RuntimeError("ModelPT runs on sm_100a only; there is no CPU path (use the reference for CPU)")
        idx = device.index if device.index is not None else torch.cuda.current_device()
        if self._device is not None and self._device.index == idx:
            return self
        self._release()
        self._device = torch.device("cuda", idx)
        self._ctx = Context.get(idx)
        ws, bs = [], []
        for k in CONV_KEYS:
            ws.append(np.ascontiguousarray(self._state[k + ".weight"].numpy(), dtype=np.float32))
            bs.append(np.ascontiguousarray(self._state[k + ".bias"].numpy(), dtype=np.float32))
        wp = (C.c_void_p * 19)(*[w.ctypes.data for w in ws])
        bp = (C.c_void_p * 19)(*[b.ctypes.data for b in bs])
        mean = (C.c_float * 3)(*self._state["net.0.mean"].flatten().tolist())
        std = (C.c_float * 3)(*self._state["net.0.std"].flatten().tolist())
        maxv = float(self._state["net.0.max_pixel_value"])
        h = C.c_void_p()
        with torch.cuda.device(idx):
            check(lib().srb_model_pt_create(self._ctx.handle, wp, bp, mean, std, maxv, LRELU_SLOPE, self._dtype,
                                            C.byref(h)), "srb_model_pt_create")
        self._net = h
        return self

    def cuda(self, device=None):
        return self.to(torch.device("cuda", device if device is not None else torch.cuda.current_device()))

    def _release(self):
        if self._net is not None:
            lib().srb_model_pt_destroy(self._net)
            self._net = None
        self._ws = None

    def __del__(self):
        try:
            self._release()
        except Exception:
            pass

    # ---- forward ---------------------------------------------------------------------------
    def _workspace(self, nbytes: int) -> torch.Tensor:
        if self._ws is None or self._ws.numel() < nbytes:
            self._ws = None
            self._ws = torch.empty(nbytes, dtype=torch.uint8, device=self._device)
        return self._ws

    def workspace_bytes(self, N: int, H: int, W: int) -> int:
        return int(lib().srb_model_pt_workspace_bytes(N, H, W))

    _IN_LAYOUTS = {"nchw": 0, "hwc_bgr": 1, "hwc_rgb": 2}

    @torch.no_grad()
    def forward(self, x: torch.Tensor, out: torch.Tensor | None = None, layout: str = "nchw", in_layout: str = "nchw") -> torch.Tensor:
        """uint8 [N,3,H,W] -> uint8 [N,3,H,W] (rs.py:64 / :94).  layout="hwc_bgr" (device inputs): the result is written as
        [N,H,W,3] with the channels reversed — the frame the video writer takes (rs.py:65-74: RGB2BGR + permute), produced
        by the tail conv's store instead of two elementwise kernels.  in_layout="hwc_rgb" / "hwc_bgr" (device inputs): x is
        [N,H,W,3] uint8 as a decoder hands frames over; the head conv's loader reads it directly (no permute, no `.float()`,
        no Normalize pass — rs.py:60-63)."""
        if layout not in ("nchw", "hwc_bgr"):
            raise ValueError(f"layout {layout!r}")
        if in_layout not in self._IN_LAYOUTS:
            raise ValueError(f"in_layout {in_layout!r}")
        if self._net is None:
            raise RuntimeError("call .to('cuda') first (as rs.py:49 does)")
        hwc_in = in_layout != "nchw"
        if x.dim() != 4 or x.shape[3 if hwc_in else 1] != 3:
            raise RuntimeError(f"expected {'[N,H,W,3]' if hwc_in else '[N,3,H,W]'}, got {tuple(x.shape)}")
        if x.dtype != torch.uint8:
            # The trace baked the uint8 branch of Normalize: a float input is divided by 255
            # too.  Only integral values in [0,255] are representable on the uint8 kernel path.
            xr = x.round()
            if not bool(((xr == x) & (x >= 0) & (x <= 255)).all()):
                raise RuntimeError("float input with non-integral / out-of-range pixels is not supported by the "
                                   "uint8 kernel path (model.pt divides by 255 internally)")
            x = xr.to(torch.uint8)
        if hwc_in:
            N, H, W, _ = x.shape
        else:
            N, _, H, W = x.shape
        oshape = (N, 3, H, W) if layout == "nchw" else (N, H, W, 3)
        if N == 0 or H == 0 or W == 0:
            return x.new_empty(oshape)
        if not x.is_cuda and (layout != "nchw" or hwc_in):
            raise RuntimeError("HWC layouts take device inputs (use stream(..., layout=, in_layout=) for host batches)")
        if not x.is_cuda:  # host tensors: the C ABI's host entry does H2D / D2H itself
            x = x.contiguous()
            y = torch.empty_like(x)
            with torch.cuda.device(self._device):
                check(lib().srb_model_pt_forward_host(self._ctx.handle, self._net, x.data_ptr(), y.data_ptr(), N, H, W),
                      "srb_model_pt_forward_host")
            return y
        if x.device != self._device:
            raise RuntimeError(f"input on {x.device}, model on {self._device}")
        x = x.contiguous()
        y = out if out is not None else x.new_empty(oshape)
        if y.numel() != x.numel() or y.dtype != torch.uint8 or not y.is_contiguous() or y.device != x.device:
            raise RuntimeError("out must be a contiguous uint8 tensor of the input's size on the input's device")
        nbytes = self.workspace_bytes(N, H, W)
        ws = self._workspace(nbytes)
        stream = C.c_void_p(torch.cuda.current_stream(self._device).cuda_stream)
        with torch.cuda.device(self._device):
            check(lib().srb_model_pt_forward_io(self._ctx.handle, self._net, x.data_ptr(), self._IN_LAYOUTS[in_layout], y.data_ptr(),
                                                1 if layout == "hwc_bgr" else 0, N, H, W, ws.data_ptr(), ws.numel(), stream),
                  "srb_model_pt_forward_io")
        return y

    __call__ = forward

    @torch.no_grad()
    def stream(self, batches, depth: int = 2, layout: str = "nchw", graphs: bool = True, in_layout: str = "nchw"):
        """Pipelined host-to-host inference over a sequence of uint8 host batches — the shape of rs.py's video loop
        (rs.py:58-76: frames from the reader -> model -> writer).  Yields one pinned uint8 host tensor per input batch, in
        order.  H2D of batch i+1 and D2H of batch i-1 run on copy streams while the kernels of batch i run on the compute
        stream; `depth` batches are in flight.  A yielded tensor stays valid until `depth` more items have been taken from the
        generator (the ring holds 2 * depth + 1 slots; the slot is then reused): write it out or copy it before that, as a
        frame writer does.  `list(model.stream(...))` of more than depth + 1 batches therefore needs a .clone() per item.
        The ring of device / pinned buffers is kept on the model between calls, and (graphs=True) the 11 launches of a slot's
        forward are captured once per (slot, shape) as a CUDA graph: a step costs the host three enqueues, which is what
        keeps eight ranks sharing one host at the single-GPU rate.  in_layout="hwc_rgb" / "hwc_bgr": the host batches are
        [N,H,W,3] frames (a decoder's layout); they are uploaded as they are and read by the head conv's loader."""
        if self._net is None:
            raise RuntimeError("call .to('cuda') first (as rs.py:49 does)")
        if layout not in ("nchw", "hwc_bgr"):
            raise ValueError(f"layout {layout!r}")
        if in_layout not in self._IN_LAYOUTS:
            raise ValueError(f"in_layout {in_layout!r}")
        hwc_in = in_layout != "nchw"
        depth = max(1, int(depth))
        dev = self._device
        st = self.__dict__.get("_stream_state")
        if st is None or st["net"] is not self._net or st["dev"] != dev:
            st = self.__dict__["_stream_state"] = dict(net=self._net, dev=dev, slots={},
                                                       streams=tuple(torch.cuda.Stream(dev) for _ in range(3)))
        s_in, s_c, s_out = st["streams"]
        nbuf = 2 * depth + 1
        pending = []                   # slots submitted and not yet yielded
        i = 0

        def make_slot(shape):
            n_, h_, w_ = (shape[0], shape[1], shape[2]) if hwc_in else (shape[0], shape[2], shape[3])
            oshape = (n_, 3, h_, w_) if layout == "nchw" else (n_, h_, w_, 3)
            x_dev = torch.empty(shape, dtype=torch.uint8, device=dev)
            y_dev = torch.empty(oshape, dtype=torch.uint8, device=dev)
            y_host = torch.empty(oshape, dtype=torch.uint8).pin_memory()
            g = None
            if graphs:
                with torch.cuda.stream(s_c):               # eager once: workspace, tensor-map caches, kernel attributes
                    self.forward(x_dev, out=y_dev, layout=layout, in_layout=in_layout)
                torch.cuda.synchronize(dev)
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=s_c, capture_error_mode="thread_local"):   # (NCCL watchdog / sampler threads may call CUDA)
                    self.forward(x_dev, out=y_dev, layout=layout, in_layout=in_layout)
            # (the captured launches hold the address of the workspace they were recorded with: the slot keeps it alive)
            return (x_dev, y_dev, y_host, torch.cuda.Event(), torch.cuda.Event(), torch.cuda.Event(), g, self._ws)

        for xb in batches:
            if xb.dtype != torch.uint8 or xb.dim() != 4 or xb.shape[3 if hwc_in else 1] != 3 or xb.is_cuda:
                raise RuntimeError("stream() takes uint8 host tensors [N,3,H,W] ([N,H,W,3] with in_layout='hwc_*')")
            xb = xb.contiguous()
            if len(pending) > depth:                       # (never: a batch is yielded as soon as depth + 1 are pending)
                raise RuntimeError("internal: ring overrun")
            key = (i % nbuf, nbuf, layout, in_layout, tuple(xb.shape), bool(graphs))
            sl = st["slots"].get(key)
            if sl is None:
                for j in pending:                          # (capturing synchronises: nothing may be in flight on a slot)
                    j[5].synchronize()
                sl = st["slots"][key] = make_slot(tuple(xb.shape))
            x_dev, y_dev, y_host, ev_in, ev_c, ev_out, g, _ = sl
            with torch.cuda.stream(s_in):
                x_dev.copy_(xb, non_blocking=True)
                ev_in.record(s_in)
            with torch.cuda.stream(s_c):
                s_c.wait_event(ev_in)
                if g is not None:
                    g.replay()
                else:
                    self.forward(x_dev, out=y_dev, layout=layout, in_layout=in_layout)
                ev_c.record(s_c)
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_c)
                y_host.copy_(y_dev, non_blocking=True)
                ev_out.record(s_out)
            pending.append(sl)
            i += 1
            if len(pending) > depth:
                j = pending.pop(0)
                j[5].synchronize()
                self._ctx.poll()                           # a kernel-side pipeline timeout raises here, not downstream
                yield j[2]
        for j in pending:
            j[5].synchronize()
            self._ctx.poll()
            yield j[2]

    @staticmethod
    def launches_per_forward(H: int = 64, W: int = 64) -> int:
        return int(lib().srb_model_pt_launches(H, W))


def load(path, map_location="cpu", dtype="fp16") -> ModelPT:
    """Same call shape as `torch.jit.load(path, "cpu")` (rs.py:48)."""
    del map_location  # weights are staged on the host until .to(device), like the reference
    return ModelPT(_load_state(path), dtype=dtype)
