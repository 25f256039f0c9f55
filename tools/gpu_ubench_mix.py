#!/usr/bin/env python
"""tcgen05.mma issue-stream micro-benchmark (probe build only: SRB200_LIB=.../libsrb200_dbg.so).

Reproduces the fused residual block's issuers on one SM: per 128-pixel row, convA = 12 MMAs of N = 96 into three adjacent
hidden rows, convB = 6 MMAs of N = 192 into three adjacent output rows, each a dependent chain on its own accumulator, and
asks which arrangement of chains the tensor pipe runs fastest (cycles per row; ideal = 12*48 + 6*96 = 1152 of math).
usage: SRB200_LIB=image_super_resolution_b200/libsrb200_dbg.so python tools/gpu_ubench_mix.py
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ.setdefault("SRB200_LIB", os.path.join(ROOT, "image_super_resolution_b200", "libsrb200_dbg.so"))
sys.path.insert(0, ROOT)
from image_super_resolution_b200._lib import Context, lib, check  # noqa: E402

ROWS = 400


def run(ctx, warps, pair=0):
    n = (C.c_int * 4)(*[w[0] for w in warps] + [16] * (4 - len(warps)))
    cnt = (C.c_int * 4)(*[w[1] for w in warps] + [0] * (4 - len(warps)))
    acc = (C.c_int * 4)(*[w[2] for w in warps] + [1] * (4 - len(warps)))
    out = (C.c_longlong * 16)()
    check(lib().srb_ubench_mix(ctx.handle, len(warps), n, cnt, acc, ROWS, pair, out), "srb_ubench_mix")
    return out[0] / ROWS, [out[1 + i] / ROWS for i in range(len(warps))], [out[5 + i] / ROWS for i in range(len(warps))]


def main():
    ctx = Context.get(0)
    cases = [
        ("convA alone: 12 x N96, one chain", [(96, 12, 1)]),
        ("convA alone: 12 x N96 over 2 accumulators (one issuer)", [(96, 12, 2)]),
        ("convA alone: 12 x N96 over 4 accumulators (one issuer)", [(96, 12, 4)]),
        ("convA as two issuers x 6 x N96 (K split, own chains)", [(96, 6, 1), (96, 6, 1)]),
        ("convB alone: 6 x N192, one chain", [(192, 6, 1)]),
        ("current kernel: issuer A 12 x N96 chain | issuer B 6 x N192 chain", [(96, 12, 1), (192, 6, 1)]),
        ("K-split: issuers A0, A1 6 x N96 chains | issuer B 6 x N192 chain", [(96, 6, 1), (96, 6, 1), (192, 6, 1)]),
        ("one issuer A over 2 accumulators | issuer B", [(96, 12, 2), (192, 6, 1)]),
        ("4 issuers: A0, A1 (6 x N96) | B0, B1 (3 x N192)", [(96, 6, 1), (96, 6, 1), (192, 3, 1), (192, 3, 1)]),
        ("single issuer, everything in order: 12 x N96 then 6 x N192 needs two calls; proxy: 18 x N128 chain", [(128, 18, 1)]),
        ("dense-block fold: one issuer, N96 over 2 accumulators (NT = 2)", [(96, 24, 2)]),
        ("dense-block fold: two issuers, N96 chains", [(96, 12, 1), (96, 12, 1)]),
        ("dense-block fold: two issuers, N96 over 2 accumulators each", [(96, 12, 2), (96, 12, 2)]),
        ("dense-block today: one issuer, N32 over 4 accumulators", [(32, 36, 4)]),
        ("dense-block: one issuer N192 chain (cout 64 fold)", [(192, 12, 1)]),
        ("dense-block: two issuers N192 chains", [(192, 6, 1), (192, 6, 1)]),
        ("N256 chain", [(256, 8, 1)]),
    ]
    print(f"cycles per row ({ROWS} rows back to back, 2 commits per row and issuer); 'issue' = issuing thread busy, 'done' = completion")
    for name, warps in cases:
        try:
            tot, iss, done = run(ctx, warps)
        except RuntimeError as e:                     # (e.g. more than 512 TMEM columns)
            print(f"{name:95s} not runnable: {str(e).split(': ')[-1]}")
            continue
        math = sum(w[0] / 2 * w[1] for w in warps)
        print(f"{name:95s} total {tot:7.0f}  math {math:5.0f}  eff {math / tot:5.2f}  issue {[round(v) for v in iss]}")
    if "--pair" in sys.argv:
        print("cta_group::2 (cluster of two CTAs, M = 256; per-SM math identical, B operand read halved):")
        for name, warps in cases[:8] + cases[10:13] + cases[14:]:
            try:
                tot, iss, done = run(ctx, warps, pair=1)
            except RuntimeError as e:
                print(f"{name:95s} FAILED {e}")
                break
            math = sum(w[0] / 2 * w[1] for w in warps)
            print(f"{name:95s} total {tot:7.0f}  math {math:5.0f}  eff {math / tot:5.2f}  issue {[round(v) for v in iss]}")
    ctx.check_device()


if __name__ == "__main__":
    main()
