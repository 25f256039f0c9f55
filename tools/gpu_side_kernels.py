#!/usr/bin/env python
"""The three non-trunk kernels of the model.pt forward (head 9x9, conv1 row-fold, tail 9x9) timed alone at the bench shape in
cold bursts, whole-row / group-major TMA boxes against one box per group (SRB_PATH_GROUP_BOXES), with a bit-equality check of
the two paths on narrow, ragged and wide shapes.   usage: python tools/gpu_side_kernels.py [--shapes-only]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import image_super_resolution_b200 as srb
from image_super_resolution_b200 import ops, _lib

dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(11)


def cold(fn, n=24, rounds=5, idle=1.0):
    v = []
    for _ in range(rounds):
        time.sleep(idle)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        fn(); fn()
        a.record()
        for _ in range(n): fn()
        b.record(); torch.cuda.synchronize()
        v.append(a.elapsed_time(b) / n * 1e3)
    v.sort()
    return v[len(v) // 2], v[0]


def make(path):
    ctx = _lib.Context(0)
    ctx.set_option(_lib.OPT_PATH, path)
    w1 = torch.randn(64, 64, 3, 3, generator=g) / 24; b1 = torch.randn(64, generator=g) * 0.1
    w2 = torch.randn(3, 64, 9, 9, generator=g) / 72; b2 = torch.randn(3, generator=g) * 0.1
    return ctx, ops.RowConv(ctx, w1, b1, "fp16"), ops.Tail9(ctx, w2, b2, "fp16")


def main():
    torch.manual_seed(3)
    shapes = [(4, 64, 64), (3, 40, 56), (5, 33, 40), (2, 24, 200), (1, 50, 136), (2, 17, 128), (1, 20, 264), (3, 9, 16), (1, 30, 1000)]
    if "--quick" in sys.argv:
        shapes = shapes[:2]
    g.manual_seed(11); ca, ra, ta = make(0)
    g.manual_seed(11); cb, rb, tb = make(_lib.PATH_GROUP_BOXES)
    ok = True
    for (N, H, W) in shapes:
        x = torch.randn(N, 8, H, W, 8, device=dev).half(); res = torch.randn(N, 8, H, W, 8, device=dev).half()
        ya = torch.empty_like(x); yb = torch.empty_like(x)
        ra.forward(x, ya, res); rb.forward(x, yb, res)
        oa = ta.forward(x * 0.2); ob = tb.forward(x * 0.2)
        fa = ta.forward(x * 0.2, u8=False); fb = tb.forward(x * 0.2, u8=False)
        torch.cuda.synchronize()
        e1 = bool((ya.view(torch.int16) == yb.view(torch.int16)).all()); e2 = bool((oa == ob).all()); e3 = bool((fa == fb).all())
        print(f"shape {N}x{H}x{W}: rowconv equal {e1}  tail u8 equal {e2}  tail f32 equal {e3}", flush=True)
        ok = ok and e1 and e2 and e3
    ca.check_device(); cb.check_device()
    print("PATHS", "BIT-EQUAL" if ok else "DIFFER", flush=True)
    if "--shapes-only" in sys.argv:
        return
    if "--quick" in sys.argv:
        x = torch.randn(256, 8, 64, 64, 8, device=dev).half(); res = torch.randn_like(x); y = torch.empty_like(x)
        out = torch.empty(256, 3, 64, 64, dtype=torch.uint8, device=dev)
        g.manual_seed(4)
        hd = ops.Head(ca, torch.randn(64, 3, 9, 9, generator=g) / 15.6, torch.randn(64, generator=g) * 0.1, "fp16")
        xi = torch.randint(0, 256, (256, 3, 64, 64), dtype=torch.uint8, device=dev); yo = torch.empty_like(xi)
        mean, std = [0.485, 0.456, 0.406], [0.229, 0.224, 0.225]
        m1 = cold(lambda: ra.forward(x, y, res)); m2 = cold(lambda: ta.forward(x, out=out)); m3 = cold(lambda: hd.forward(xi, mean, std, 0.2, out=y))
        m4 = cold(lambda: ops.c1tail_forward(ra, ta, x, res, out=out))
        print(f"QUICK conv1 + tail fused {m4[0]:6.1f} us (min {m4[1]:.1f})   two launches {m1[0] + m2[0]:6.1f} us", flush=True)
        model = srb.load(os.path.join(ROOT, "tests", "golden", "model_pt_weights.npz")).to(dev).eval()
        for _ in range(3): model(xi, out=yo)
        f = cold(lambda: model(xi, out=yo), 12)
        print(f"QUICK {os.environ.get('SRB200_LIB', 'product')[-16:]:16s} rowconv {m1[0]:6.1f}  tail {m2[0]:6.1f}  head {m3[0]:6.1f}  forward {f[0]:6.1f} us = {256 * 64 * 64 / f[0]:.0f} Mpix/s", flush=True)
        ca.check_device()
        return
    x = torch.randn(256, 8, 64, 64, 8, device=dev).half(); res = torch.randn_like(x); y = torch.empty_like(x)
    out = torch.empty(256, 3, 64, 64, dtype=torch.uint8, device=dev)
    xw = torch.randn(2, 8, 1080, 1920, 8, device=dev).half(); yw = torch.empty_like(xw)
    outw = torch.empty(2, 3, 1080, 1920, dtype=torch.uint8, device=dev)
    for rnd in range(2):
        for name, r, t in (("rows", ra, ta), ("groups", rb, tb)):
            m1 = cold(lambda: r.forward(x, y, res)); m2 = cold(lambda: t.forward(x, out=out))
            m3 = cold(lambda: r.forward(xw, yw, xw), n=6, rounds=3); m4 = cold(lambda: t.forward(xw, out=outw), n=6, rounds=3)
            print(f"round {rnd} {name:7s} 256x64x64: rowconv {m1[0]:6.1f} us (min {m1[1]:.1f})  tail {m2[0]:6.1f} us (min {m2[1]:.1f})"
                  f"   2x1080p: rowconv {m3[0]:7.1f}  tail {m4[0]:7.1f}", flush=True)
    # the ingest head alone, and the whole forward with it against the former Normalize pass + pixel-pair conv0
    g.manual_seed(4)
    hd = ops.Head(ca, torch.randn(64, 3, 9, 9, generator=g) / 15.6, torch.randn(64, generator=g) * 0.1, "fp16")
    xi = torch.randint(0, 256, (256, 3, 64, 64), dtype=torch.uint8, device=dev); yo = torch.empty_like(xi)
    xh = xi.permute(0, 2, 3, 1).contiguous()
    mean, std = [0.485, 0.456, 0.406], [0.229, 0.224, 0.225]
    m1 = cold(lambda: hd.forward(xi, mean, std, 0.2, out=y)); m2 = cold(lambda: hd.forward(xh, mean, std, 0.2, layout="hwc_bgr", out=y))
    xiw = torch.randint(0, 256, (2, 3, 1080, 1920), dtype=torch.uint8, device=dev)
    m3 = cold(lambda: hd.forward(xiw, mean, std, 0.2, out=yw), n=6, rounds=3)
    print(f"ingest head 256x64x64: nchw {m1[0]:.1f} us (min {m1[1]:.1f})  hwc_bgr {m2[0]:.1f} us   2x1080p nchw {m3[0]:.1f} us", flush=True)
    model = srb.load(os.path.join(ROOT, "tests", "golden", "model_pt_weights.npz")).to(dev).eval()
    for _ in range(3): model(xi, out=yo)
    for rnd in range(2):
        for name, path in (("ingest", 0), ("pair4+pack", _lib.PATH_PAIR4_HEAD)):
            model._ctx.set_option(_lib.OPT_PATH, path)
            for _ in range(2): model(xi, out=yo)
            f = cold(lambda: model(xi, out=yo), 12)
            print(f"round {rnd} forward 256x64x64 [{name:18s}]: {f[0]:.1f} us (min {f[1]:.1f}) = {256 * 64 * 64 / f[0]:.0f} Mpix/s", flush=True)
    model._ctx.set_option(_lib.OPT_PATH, 0)
    model._ctx.check_device()


if __name__ == "__main__":
    main()
