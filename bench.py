#!/usr/bin/env python
"""bench.py — the headline metric of BASELINE.json: output Mpix/s of model.pt inference.

    python bench.py --gpus N --steps K --warmup W            # this repo's sm_100a path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (oracle port)

Workload (config.workload): BASELINE.json configs[1] — a [256,3,64,64] uint8 batch through the model.pt
graph, fp16 operands / fp32 accumulate, on each of N GPUs (frames are independent: weak scaling, no
collective).  One step = one forward over that batch.  The working set of a step
(~487 MB of C8-plane activations) is far larger than the 126 MB L2, so no explicit L2 flush is needed.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WEIGHTS = os.path.join(ROOT, "tests", "golden", "model_pt_weights.npz")
BATCH, HH, WW = 256, 64, 64
FLOP_PER_PX = 725760  # algorithmic FLOP (2*MAC) per pixel of the model.pt graph
# algorithmic FLOP/px of each launch of one forward, in launch order (conv0 with the uint8 ingest folded in, 8x(A,B), conv1, conv2)
LAUNCH_FLOP_PX = [2 * 243 * 64] + [2 * 576 * 32, 2 * 288 * 64] * 8 + [2 * 576 * 64, 2 * 5184 * 3]
LAUNCH_NAMES = ["ingest_conv9x9_head_3_64"] + ["conv3x3_64_32", "conv3x3_32_64"] * 8 + ["conv3x3_64_64", "conv9x9_tail_64_3"]
assert sum(LAUNCH_FLOP_PX) == FLOP_PER_PX
# with the fused residual-block kernel (widths that qualify, srb_resblock_supported): 11 launches, 10 when conv1 and the tail
# run as one launch (rows of at least 57 pixels, srb_c1tail_supported)
FUSED_FLOP_PX = [2 * 243 * 64] + [2 * 576 * 32 + 2 * 288 * 64] * 8 + [2 * 576 * 64, 2 * 5184 * 3]
FUSED_NAMES = ["ingest_conv9x9_head_3_64"] + ["resblock_64_32_64_fused"] * 8 + ["conv3x3_64_64", "conv9x9_tail_64_3"]
FUSED10_FLOP_PX = FUSED_FLOP_PX[:9] + [2 * 576 * 64 + 2 * 5184 * 3]
FUSED10_NAMES = FUSED_NAMES[:9] + ["conv3x3_64_64_skip_conv9x9_tail_64_3_fused"]
assert sum(FUSED10_FLOP_PX) == FLOP_PER_PX
assert sum(FUSED_FLOP_PX) == FLOP_PER_PX
# DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) of the dominant kernel at the bench shape, from the
# one `ncu --set full` capture of a whole forward summarised in profiles/ncu_fwd_r1_full.txt (mean of the 8 launches:
# 143.8 MB read + 92.85 MB written; the algorithmic traffic is 2 x 134.2 MB, part of the output is still in the 126 MB
# L2 when the launch ends).
NCU_TRAFFIC_BYTES = {"resblock_64_32_64_fused": 236.65e6}
NCU_TRAFFIC_SHAPE = (256, 64, 64)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"bf16_tflops": d.get("bf16_tflops"), "bf16_tflops_sustained": d.get("bf16_tflops_sustained"),
                "hbm_gbs": d.get("hbm_gbs"), "source": "measured"}
    return {"bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "hbm_gbs": 6650.0, "source": "fallback"}


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index: int, period: float = 0.004):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.power, self.reasons, self.max_mhz = [], [], set(), None
        self._halt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # noqa: BLE001
            self.err = repr(e)

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
                 nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake"}
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                self.power.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(self.period)

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        pw = sorted(self.power)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s), "power_w_median": (pw[len(pw) // 2] if pw else None), "power_w_max": (pw[-1] if pw else None)}


def dist_setup(n_gpus: int):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def cpu_port_forward_timer(sample_tiles: int, min_seconds: float, max_iters: int = 1000):
    """Times the oracle (CPU fp32 port of the reference path, all host threads) on a bounded sample."""
    from oracle import modelpt_oracle as mo  # the one place bench.py executes oracle/: the CPU baseline
    torch.set_num_threads(os.cpu_count() or 1)
    sd = mo.load_weights(WEIGHTS)
    g = torch.Generator().manual_seed(100)
    x = torch.randint(0, 256, (sample_tiles, 3, HH, WW), dtype=torch.uint8, generator=g)
    mo.forward_u8(x, sd)  # warm-up
    times = []
    t_all = time.perf_counter()
    while (time.perf_counter() - t_all < min_seconds or len(times) < 3) and len(times) < max_iters:
        t0 = time.perf_counter()
        mo.forward_u8(x, sd)
        times.append(time.perf_counter() - t0)
    px = sample_tiles * HH * WW
    return px, times


def run_reference(args, rank, world):
    """Reference arm: the reference's own CPU implementation of the path (fp32, oneDNN through ATen, all host
    threads), restated by oracle/ from the extracted weights — /root/reference does not exist on the GPU box."""
    if rank != 0:
        return
    sample_tiles = 32  # bounded sample of the [256,3,64,64] workload per step
    from oracle import modelpt_oracle as mo
    torch.set_num_threads(os.cpu_count() or 1)
    sd = mo.load_weights(WEIGHTS)
    g = torch.Generator().manual_seed(100)
    x = torch.randint(0, 256, (sample_tiles, 3, HH, WW), dtype=torch.uint8, generator=g)
    for _ in range(max(args.warmup, 1)):
        mo.forward_u8(x, sd)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        mo.forward_u8(x, sd)
    dt = time.perf_counter() - t0
    px = sample_tiles * HH * WW * args.steps
    v = px / dt / 1e6
    sample = f"{sample_tiles} of the 256 64x64 tiles per step, fp32, torch CPU ({torch.get_num_threads()} threads)"
    line = {"impl": "reference", "metric": "model.pt inference output Mpix/s", "value": v, "unit": "Mpix/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": max(args.warmup, 1), "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "model.pt batched inference, 64x64 LR tiles x256 batch (BASELINE configs[1]); "
                                   "reference arm runs a bounded 32-tile sample per step on host cores"},
            "cpu_baseline": {"value": v, "unit": "Mpix/s", "cores": torch.get_num_threads(), "kind": "port",
                             "sample": sample},
            "e2e": {"value": v, "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def time_cuda(fn, steps, warmup):
    for _ in range(warmup):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def gpu_reference(dev, model, steps, warmup):
    """The number the kernels must beat: the same model.pt graph as STOCK PyTorch CUDA ops —
    aten::_convolution -> cuDNN, leaky_relu_, add, tanh, round — under torch.autocast('cuda') (fp16) with cudnn.benchmark and
    the TF32 flags, exactly as the reference runs it on a GPU (rs.py:62-64, :93-94, :126-129), on the same B200, same inputs,
    same steps / warm-up, CUDA-event timed.  A baseline leg: it executes the oracle's restatement of the graph on device
    tensors; none of this repo's kernels are on it."""
    from oracle import modelpt_oracle as mo
    torch.backends.cudnn.benchmark = True
    torch.backends.cuda.matmul.allow_tf32 = True
    torch.backends.cudnn.allow_tf32 = True
    sd = {k: v.to(dev) for k, v in mo.load_weights(WEIGHTS).items()}
    pk = peaks()
    out = {"what": "stock PyTorch (cuDNN) model.pt graph, torch.autocast fp16, cudnn.benchmark, TF32 flags (rs.py:62-64,126-129)",
           "torch": torch.__version__, "cudnn": torch.backends.cudnn.version()}
    g = torch.Generator().manual_seed(100)
    for name, shape in (("config2_256x64x64", (BATCH, 3, HH, WW)), ("config3_2x1080p", (2, 3, 1080, 1920))):
        x = torch.randint(0, 256, shape, dtype=torch.uint8, generator=g).to(dev)

        def ref_fwd():
            with torch.autocast(device_type="cuda"):
                return mo.forward_u8(x, sd)

        y = torch.empty_like(x)
        px = shape[0] * shape[2] * shape[3]
        ms_ref = time_cuda(ref_fwd, steps, max(warmup, 5))       # (cudnn.benchmark picks its algorithms during warm-up)
        ms_ours = time_cuda(lambda: model(x, out=y), steps, max(warmup, 3))
        d = (ref_fwd().int() - y.int()).abs()
        out[name] = {"reference_mpix_s": px / ms_ref / 1e3, "reference_ms": ms_ref,
                     "reference_frac_of_bf16_peak": px * FLOP_PER_PX / (ms_ref * 1e-3) / 1e12 / pk["bf16_tflops"],
                     "ours_mpix_s": px / ms_ours / 1e3, "ours_ms": ms_ours, "ours_over_gpu_reference": ms_ref / ms_ours,
                     "max_abs_lsb_vs_ours": int(d.max()), "frac_exact_vs_ours": float((d == 0).float().mean())}
        del x, y
        torch.cuda.empty_cache()
    return out


def sustained(dev, local, model, x, y, seconds=3.0):
    """>= `seconds` of back-to-back forwards (the headline `value` is an 18 ms burst): the rate a video stream sees once the
    1000 W power cap has engaged, with NVML clock / power samples, against the SUSTAINED cuBLAS bf16 figure."""
    px_step = x.shape[0] * x.shape[2] * x.shape[3]
    for _ in range(3):
        model(x, out=y)
    torch.cuda.synchronize()
    sampler = ClockSampler(local, period=0.02)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.start()
    t0 = time.perf_counter()
    e0.record()
    n = 0
    while True:
        for _ in range(100):
            model(x, out=y)
        n += 100
        if time.perf_counter() - t0 > seconds:     # (the host runs ahead of the device: the timed region is >= this long)
            break
    e1.record()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    ms = e0.elapsed_time(e1)
    pk = peaks()
    tf = n * px_step * FLOP_PER_PX / (ms * 1e-3) / 1e12
    return {"value": n * px_step / ms / 1e3, "unit": "Mpix/s", "seconds": ms / 1e3, "steps": n, "ms_per_step": ms / n,
            "tflops": tf, "frac_of_bf16_sustained": tf / pk["bf16_tflops_sustained"], "frac_of_bf16_burst": tf / pk["bf16_tflops"],
            "peak_sustained": pk["bf16_tflops_sustained"], "clocks": clocks}


class _StubBoard:
    """The reference always logs the loss per step (train.py:64-65: loss.item() + tensorBoard.add_scalar)."""

    def __init__(self):
        self.n = 0

    def add_scalar(self, *a, **k):
        self.n += 1


def bench_train(dev, rank, world, steps, warmup=3):
    """BASELINE configs[4]: train.py step (EResNet(16,0.2,2), L1 loss, GradScaler, clip, Adam, LinearLR, EMA) on synthetic
    COCO-shaped 128x128 HR crops, per-GPU batch 16, data parallel with one gradient all-reduce per step.  Runs the
    drop-in `train()` loop itself; samples/s = global batch x steps / max-over-ranks device time."""
    import torch.distributed as dist
    from image_super_resolution_b200 import models
    from image_super_resolution_b200 import train as T
    from image_super_resolution_b200.training import L1Loss
    B = 16
    g = torch.Generator().manual_seed(100 + rank)
    hr = (torch.rand(B, 3, 128, 128, generator=g) * 2 - 1).to(dev)
    lr = torch.randn(B, 3, 64, 64, generator=g).to(dev)
    loss_fn = L1Loss().to(dev)

    def one(tail: str, board):
        """ms per step (max over ranks) of the drop-in train() with the given optimizer tail and logger."""
        torch.manual_seed(100)
        net = models.EResNet(16, 0.2, 2).to(dev)
        T.broadcast_parameters(net)
        # the reference's optimizer (train.py:265: Adam, lr 1e-4, betas (0.9, 0.999), no weight decay), GradScaler, gradient
        # clipping and EMA: either the library's fused tail (optim.py: two launches over flat arenas, same arithmetic and call
        # order; parity vs the torch objects in tests/test_gpu_optim.py) or torch's own objects, unmodified
        if tail == "torch":
            opt = torch.optim.Adam(net.parameters(), lr=1e-4, betas=(0.9, 0.999))
            scaler, ema = torch.amp.GradScaler("cuda"), T.ModelEMA(net, tau=1000)
        else:
            opt = T.FlatAdam(net.parameters(), lr=1e-4, betas=(0.9, 0.999))
            scaler, ema = T.GradScaler("cuda"), T.ModelEMA(net, tau=1000).attach(opt)
        sched = torch.optim.lr_scheduler.LinearLR(opt, start_factor=1, end_factor=0.1, total_iters=1000)
        T.train(net, ema, [(hr, lr)] * warmup, loss_fn, opt, scaler, sched, 0, board)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0.record()
        losses = T.train(net, ema, [(hr, lr)] * steps, loss_fn, opt, scaler, sched, 1, board)
        e1.record()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        del net, opt, ema
        return float(ms.item()) / steps, losses

    # headline of the leg: the reference's loop as the reference runs it — loss.item() + add_scalar EVERY step (train.py:64-65)
    ms, losses = one("fused", _StubBoard())
    ms_torch, _ = one("torch", _StubBoard())
    ms_deferred, _ = one("fused", None)          # no logger attached: train() takes the read-backs together at the end
    flop = 289.0e9 * B          # fwd + bwd algorithmic FLOP per sample
    pk = peaks()
    return {"metric": "train.py step samples/s", "value": world * B / ms * 1e3, "unit": "samples/s", "ms_per_step": ms,
            "steps": steps, "model": "EResNet(16,0.2,2)", "per_gpu_batch": B, "hr_crop": 128, "operands": "fp16 in + GradScaler loss scaling (the reference's autocast precision), fp32 accumulate",
            "loop": "drop-in train(): autocast fwd, L1, GradScaler backward, unscale, clip(10), Adam, LinearLR, EMA, per-step "
                    "loss.item() + add_scalar (train.py:64-65); fused tail (FlatAdam + GradScaler + clip + EMA: 2 launches)",
            "ms_per_step_torch_tail": ms_torch, "samples_s_torch_tail": world * B / ms_torch * 1e3,
            "torch_tail": "torch.amp.GradScaler / torch clip_grad_norm_ / torch.optim.Adam / per-tensor EMA on the same fwd+bwd kernels",
            "ms_per_step_no_logger": ms_deferred,
            "collective": "one all-reduce (mean) of the 11.7M-element fp32 gradient arena per step" if world > 1 else "none",
            "fwd_bwd_tflops_per_gpu": flop / (ms * 1e-3) / 1e12, "frac_of_bf16_peak": flop / (ms * 1e-3) / 1e12 / pk["bf16_tflops"],
            "loss_first_last": [losses[0], losses[-1]]}


def run_ours(args, rank, world, local):
    import torch.distributed as dist
    import image_super_resolution_b200 as srb
    from image_super_resolution_b200._lib import lib, check

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # stdout carries exactly one JSON line: NCCL prints its version banner to stdout when the communicator comes up
        # (eagerly here, with device_id) — send file descriptor 1 to stderr while that happens
        sys.stdout.flush()
        saved_fd = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_fd, 1)
            os.close(saved_fd)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    model = srb.load(WEIGHTS).to(dev).eval()
    g = torch.Generator().manual_seed(100 + rank)
    x_host = torch.randint(0, 256, (BATCH, 3, HH, WW), dtype=torch.uint8, generator=g).pin_memory()
    y_host = torch.empty_like(x_host).pin_memory()
    x = x_host.to(dev)
    y = torch.empty_like(x)
    K, Wm = args.steps, max(args.warmup, 3)
    px_step = BATCH * HH * WW

    for _ in range(Wm):
        model(x, out=y)
    # ---- device-resident throughput ("value") ----
    sampler = ClockSampler(local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    sampler.start()
    e0.record()
    for _ in range(K):
        model(x, out=y)
    e1.record()
    barrier()
    clocks = sampler.stop()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())

    # ---- end to end through the public API, host buffers, copies inside the timed region ----
    # `ModelPT.stream` is the public call for a sequence of host batches (the shape of rs.py's video loop): every step's
    # input is copied from pinned host memory and its result copied back to pinned host memory inside the timed region;
    # the copies of neighbouring steps overlap the kernels of the current one (two batches in flight).
    def host_batches(n):
        for _ in range(n):
            yield x_host
    depth = int(os.environ.get("SRB_STREAM_DEPTH", "2"))
    for yh in model.stream(host_batches(3 + depth), depth=depth):
        pass
    barrier()
    e0.record()
    n_out = 0
    for yh in model.stream(host_batches(K), depth=depth):
        n_out += 1
    y_host.copy_(yh)
    e1.record()
    assert n_out == K
    barrier()
    ms2 = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms2, op=dist.ReduceOp.MAX)
    ms_e2e = float(ms2.item())

    # ---- per-launch durations (CUDA events recorded by the library on the launching stream) ----
    ctxh = model._ctx.handle
    nl = model.launches_per_forward(HH, WW)
    names, flops = ((FUSED10_NAMES, FUSED10_FLOP_PX) if nl == len(FUSED10_NAMES) else (FUSED_NAMES, FUSED_FLOP_PX) if nl == len(FUSED_NAMES)
                    else (LAUNCH_NAMES, LAUNCH_FLOP_PX))
    acc = [0.0] * nl
    buf = (C.c_float * 64)()
    check(lib().srb_profile_enable(ctxh, 1))
    for _ in range(K):
        model(x, out=y)
        n = lib().srb_profile_read(ctxh, buf, 64)
        assert n == nl, (n, nl)
        for i in range(nl):
            acc[i] += buf[i]
    # the trunk's eight fused blocks overlap through programmatic dependent launch (a block's prologue runs under its predecessor's
    # tail): their AVERAGE duration inside the pipeline = the trunk's span / 8, events only around the trunk
    trunk_ms = None
    if nl in (len(FUSED_NAMES), len(FUSED10_NAMES)):
        check(lib().srb_profile_enable(ctxh, 2))
        t_acc = 0.0
        for _ in range(K):
            model(x, out=y)
            n = lib().srb_profile_read(ctxh, buf, 64)
            assert n == 3, n
            t_acc += buf[1]
        trunk_ms = t_acc / K
    check(lib().srb_profile_enable(ctxh, 0))
    per_launch = [a / K for a in acc]
    groups = {}
    for name, t, f in zip(names, per_launch, flops):
        gsum = groups.setdefault(name, [0, 0.0, 0.0])
        gsum[0] += 1; gsum[1] += t; gsum[2] += f * px_step
    pk = peaks()
    kernels = {k: {"launches": v[0], "ms_per_launch": v[1] / v[0], "tflops": (v[2] / v[0]) / (v[1] / v[0]) / 1e9,
                   "share": v[1] / sum(per_launch)} for k, v in groups.items()}
    top = max(groups, key=lambda k: groups[k][1])
    isolated = kernels[top]["tflops"]
    achieved = isolated
    pipe = None
    if trunk_ms is not None and top == "resblock_64_32_64_fused":
        # average launch duration of the dominant kernel as it runs in the timed region (overlapping launches)
        ms_launch = trunk_ms / 8
        achieved = (groups[top][2] / groups[top][0]) / ms_launch / 1e9
        pipe = {"ms_per_launch": ms_launch, "tflops": achieved, "how": "CUDA events before the first and after the last of the eight "
                "back-to-back launches, / 8", "isolated_ms_per_launch": kernels[top]["ms_per_launch"], "isolated_tflops": isolated,
                "isolated_frac": isolated / pk["bf16_tflops"],
                "isolated_how": "an event after every launch (ends the overlap between launches)"}
    roofline = {"bound": "tensor", "kernel": top, "achieved": achieved, "peak": pk["bf16_tflops"], "unit": "TFLOP/s",
                "frac": achieved / pk["bf16_tflops"], "in_pipeline": pipe,
                "traffic": NCU_TRAFFIC_BYTES.get(top) if (BATCH, HH, WW) == NCU_TRAFFIC_SHAPE else None,
                "peak_source": f"{pk['source']} burst bf16 (MEASURED_PEAKS.json)",
                "whole_step_tflops": px_step * FLOP_PER_PX / (sum(per_launch) / 1e3) / 1e12,
                "kernels": kernels}

    gpu_ref = sus = None
    if world == 1 and not args.no_gpu_reference:
        try:
            gpu_ref = gpu_reference(dev, model, K, Wm)
        except Exception as e:  # noqa: BLE001
            gpu_ref = {"error": repr(e)[:300]}
    if not args.no_sustained:
        try:
            sus = sustained(dev, local, model, x, y, seconds=args.sustained_seconds)
        except Exception as e:  # noqa: BLE001
            sus = {"error": repr(e)[:300]}

    train_line = None
    if not args.no_train:
        try:
            train_line = bench_train(dev, rank, world, steps=min(K, 10))
        except Exception as e:  # noqa: BLE001  (the headline line must still print)
            train_line = {"error": repr(e)[:300]}

    if rank == 0:
        value = world * px_step * K / (ms_total / 1e3) / 1e6
        e2e = world * px_step * K / (ms_e2e / 1e3) / 1e6
        line = {"metric": "model.pt inference output Mpix/s", "value": value, "unit": "Mpix/s", "n_gpus": world,
                "steps": K, "warmup": Wm, "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f16", "data": "synthetic",
                "config": {"workload": "model.pt batched inference, 64x64 LR tiles x256 batch per GPU (BASELINE configs[1])",
                           "per_gpu_batch": BATCH, "tile": [HH, WW], "operands": "fp16 in, fp32 accumulate",
                           "l2": "working set 487 MB per step > 126 MB L2 (no flush needed)",
                           "parallelism": f"frames sharded over {world} GPU(s), no collective"},
                "tflops": value * 1e6 * FLOP_PER_PX / 1e12,
                "frac_of_bf16_peak": value * 1e6 * FLOP_PER_PX / 1e12 / (pk["bf16_tflops"] * world),
                "clocks": clocks,
                "e2e": {"value": e2e, "unit": "Mpix/s", "h2d_bytes_per_step": x_host.numel(),
                        "d2h_bytes_per_step": y_host.numel(), "ms_per_step": ms_e2e / K,
                        "api": f"ModelPT.stream(host batches): pinned H2D + forward + D2H per step, {depth} batches in flight"},
                "gpu_launches": nl * K,
                "roofline": roofline}
        if gpu_ref is not None:
            line["gpu_reference"] = gpu_ref
        if sus is not None:
            line["sustained"] = sus
        if train_line is not None:
            line["train_step"] = train_line
        if world == 1 and not args.no_cpu:
            px, times = cpu_port_forward_timer(sample_tiles=32, min_seconds=10.0)
            best = min(times)
            line["cpu_baseline"] = {"value": px / best / 1e6, "unit": "Mpix/s", "cores": torch.get_num_threads(),
                                    "kind": "port",
                                    "sample": f"32 of the 256 64x64 tiles, fp32 torch CPU oracle, best of {len(times)} runs"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-train", action="store_true", help="skip the train_step leg (BASELINE configs[4])")
    ap.add_argument("--no-gpu-reference", action="store_true", help="skip the stock-PyTorch (cuDNN fp16) leg on the same GPU")
    ap.add_argument("--no-sustained", action="store_true", help="skip the >= 3 s back-to-back leg")
    ap.add_argument("--sustained-seconds", type=float, default=3.0)
    args = ap.parse_args()
    rank, world, local = dist_setup(args.gpus)
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local)


if __name__ == "__main__":
    main()
