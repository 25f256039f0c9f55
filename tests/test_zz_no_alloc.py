"""Allocation-free entry point srb_l1_loss_fwd_noalloc (include/srb200.h; no counterpart in the reference):
its reduction scratch is a file-scope `static __device__` array of the library instead of a
caller-provided buffer.  Called through the C ABI and compared with the existing entry point srb_l1_loss_fwd (whose
`partials` the caller allocates) at the tolerance the L1 tests use (rel 1e-6), on two calls in a row that share the array.

Collected last (file name).  Passes alone in a fresh process and after the other tests.
"""
import ctypes as C
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

MAX_ELEMS = 67108864        # SRB_L1_NOALLOC_MAX_ELEMS


def _both(lib, ctx, pred, tgt, w):
    """(existing entry point with caller-provided partials, allocation-free entry point) on the current stream."""
    from image_super_resolution_b200._lib import check
    dev = pred.device
    code = {torch.float16: 0, torch.bfloat16: 1, torch.float32: 2}[pred.dtype]
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    ref = torch.full((), -1.0, dtype=torch.float32, device=dev)
    out = torch.full((), -2.0, dtype=torch.float32, device=dev)
    partials = torch.empty(lib.srb_l1_partials_len(), dtype=torch.float32, device=dev)
    check(lib.srb_l1_loss_fwd(ctx.handle, pred.data_ptr(), code, tgt.data_ptr(), pred.numel(), float(w), ref.data_ptr(),
                              partials.data_ptr(), st), "srb_l1_loss_fwd")
    free0 = torch.cuda.mem_get_info(dev)[0]
    check(lib.srb_l1_loss_fwd_noalloc(ctx.handle, pred.data_ptr(), code, tgt.data_ptr(), pred.numel(), float(w), out.data_ptr(), st),
          "srb_l1_loss_fwd_noalloc")
    torch.cuda.synchronize(dev)
    free1 = torch.cuda.mem_get_info(dev)[0]
    return float(ref.item()), float(out.item()), free0, free1


def test_l1_loss_noalloc_matches_existing_entry_point_twice_in_a_row():
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    from image_super_resolution_b200._lib import Context, lib
    dev = torch.device("cuda:0")
    ctx, l = Context.get(0), lib()
    g = torch.Generator().manual_seed(31)
    # first call: many chunks (49 CTAs), fp32 predictions, a ragged tail (n % 4 != 0); second call, right behind it on the same
    # stream: fewer chunks (stale partial sums of the first call stay in the array beyond its range), fp16 predictions
    cases = [((16, 3, 128, 128 + 1), torch.float32, 1.0), ((2, 3, 40, 56), torch.float16, 0.5), ((1, 3, 7, 5), torch.bfloat16, 2.0)]
    for rnd in range(2):                                      # ... and the whole sequence again (the array keeps its contents)
        for shape, dt, w in cases:
            pred = (torch.rand(shape, generator=g) * 2 - 1).to(dt).to(dev)
            tgt = (torch.rand(shape, generator=g) * 2 - 1).to(dev)
            ref, out, free0, free1 = _both(l, ctx, pred, tgt, w)
            want = float((pred.float() - tgt).abs().double().mean().item()) * w
            assert abs(out - ref) <= 1e-6 * abs(ref), (rnd, shape, out, ref)
            assert abs(out - want) <= 1e-5 * abs(want), (rnd, shape, out, want)
            if rnd == 1:                                      # (round 0 may load the kernel's module lazily)
                assert free1 == free0, "the allocation-free entry point changed the device's free memory"
    ctx.check_device()


def test_l1_loss_noalloc_rejects_sizes_beyond_its_scratch():
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    from image_super_resolution_b200._lib import Context, lib
    dev = torch.device("cuda:0")
    ctx, l = Context.get(0), lib()
    small = torch.zeros(64, dtype=torch.float32, device=dev)
    out = torch.zeros((), dtype=torch.float32, device=dev)
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    # (the size is checked before anything is read: no 256 MB tensors needed)
    rc = l.srb_l1_loss_fwd_noalloc(ctx.handle, small.data_ptr(), 2, small.data_ptr(), MAX_ELEMS + 1, 1.0, out.data_ptr(), st)
    assert rc != 0 and b"exceed" in l.srb_last_error()
    assert l.srb_l1_loss_fwd_noalloc(ctx.handle, small.data_ptr(), 2, small.data_ptr(), 0, 1.0, out.data_ptr(), st) != 0
    assert l.srb_l1_loss_fwd_noalloc(ctx.handle, small.data_ptr(), 2, small.data_ptr(), 64, 1.0, out.data_ptr(), st) == 0
    torch.cuda.synchronize(dev)
    assert float(out.item()) == 0.0
