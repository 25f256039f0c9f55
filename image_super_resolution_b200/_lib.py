"""ctypes binding of libsrb200.so (C ABI in include/srb200.h).

There is NO CPU or PyTorch fallback: if the shared library is missing or a call fails, this module
raises.  PyTorch is used by the callers only for device memory, streams and torch.distributed.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SRB200_LIB") or os.path.join(_HERE, "libsrb200.so")   # override: kernel-variant sweeps (tools/)

SRB_DT_F16, SRB_DT_BF16, SRB_DT_F32 = 0, 1, 2
F_LRELU, F_RESIDUAL, F_TANH, F_U8OUT, F_PIXSHUF2, F_F32OUT, F_MASK = 1, 2, 4, 8, 16, 32, 64

_vp, _i, _f, _sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
_fp = C.POINTER(C.c_float)

# name -> (restype, argtypes); every symbol declared in include/srb200.h
SIGNATURES = {
    "srb_last_error": (C.c_char_p, []),
    "srb_version": (_i, []),
    "srb_create": (_i, [_i, C.POINTER(_vp)]),
    "srb_destroy": (None, [_vp]),
    "srb_num_sms": (_i, [_vp]),
    "srb_set_option": (_i, [_vp, _i, _i]),
    "srb_pack_conv": (_i, [_vp, _vp, _vp, _vp, _f, _i, _i, _i, _i, C.POINTER(_vp)]),
    "srb_free_weights": (None, [_vp]),
    "srb_conv2d_fwd": (_i, [_vp, _vp, _vp, _i, _i, _vp, _i, _i, _vp, _i, _i, _f, _f, _f, _i, _i, _i, _i, _vp]),
    "srb_conv2d_fwd_res2": (_i, [_vp, _vp, _vp, _i, _i, _vp, _i, _i, _vp, _i, _i, _f, _vp, _i, _i, _f, _f, _f, _i, _i, _i, _i, _vp]),
    "srb_u8nchw_to_c8": (_i, [_vp, _vp, _vp, _fp, _fp, _f, _i, _i, _i, _i, _vp]),
    "srb_f32nchw_to_c8": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "srb_c8_to_f32nchw": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "srb_model_pt_create": (_i, [_vp, C.POINTER(_vp), C.POINTER(_vp), _fp, _fp, _f, _f, _i, C.POINTER(_vp)]),
    "srb_model_pt_destroy": (None, [_vp]),
    "srb_model_pt_workspace_bytes": (_sz, [_i, _i, _i]),
    "srb_model_pt_forward": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _vp, _sz, _vp]),
    "srb_model_pt_forward_frames": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _sz, _vp]),
    "srb_model_pt_forward_io": (_i, [_vp, _vp, _vp, _i, _vp, _i, _i, _i, _i, _vp, _sz, _vp]),
    "srb_c1tail_supported": (_i, [_i, _i]),
    "srb_c1tail_fwd": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "srb_head_create": (_i, [_vp, _vp, _vp, _i, C.POINTER(_vp)]),
    "srb_head_destroy": (None, [_vp]),
    "srb_head_fwd": (_i, [_vp, _vp, _vp, _i, _fp, _fp, C.c_float, _vp, C.c_float, _i, _i, _i, _vp]),
    "srb_model_pt_forward_host": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i]),
    "srb_copy_rect_u8": (_i, [_vp, _vp, _sz, _sz, _vp, _sz, _sz, _i, _sz, _sz, _i, _vp]),
    "srb_model_pt_launches": (_i, [_i, _i]),
    "srb_check_device": (_i, [_vp]),
    "srb_poll_abort": (_i, [_vp]),
    "srb_profile_enable": (_i, [_vp, _i]),
    "srb_profile_read": (_i, [_vp, _fp, _i]),
    "srb_conv2d_wgrad": (_i, [_vp, _vp, _i, _i, _i, _vp, _i, _i, _i, _vp, _f, _i, _i, _i, _i, _i, _vp]),
    "srb_bias_grad": (_i, [_vp, _vp, _i, _i, _i, _vp, _f, _i, _i, _i, _i, _vp]),
    "srb_conv2d_ex": (_i, [_vp, _vp, _vp, _vp]),
    "srb_conv_chain_create": (_i, [_vp, _i, C.POINTER(_vp), _vp, C.POINTER(_vp)]),
    "srb_conv_chain_run": (_i, [_vp, _vp, _vp]),
    "srb_conv_chain_destroy": (None, [_vp]),
    "srb_alloc_conv": (_i, [_vp, _i, _i, _i, _i, C.POINTER(_vp)]),
    "srb_alloc_conv_ex": (_i, [_vp, _i, _i, _i, _i, _i, C.POINTER(_vp)]),
    "srb_pack_conv_device": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp, _vp]),
    "srb_conv9_wgrad": (_i, [_vp, _vp, _i, _i, _i, _vp, _i, _i, _i, _vp, _f, _i, _i, _i, _i, _vp]),
    "srb_tanh_bwd_to_c8": (_i, [_vp, _vp, _vp, _vp, _f, _i, _i, _i, _i, _vp]),
    "srb_pixunshuffle_mask": (_i, [_vp, _vp, _vp, _vp, _i, _f, _i, _i, _i, _i, _vp]),
    "srb_pixshuffle_lrelu": (_i, [_vp, _vp, _vp, _i, _f, _i, _i, _i, _i, _vp]),
    "srb_c8_axpby_mask": (_i, [_vp, _vp, _i, _i, _f, _vp, _i, _i, _f, _vp, _i, _i, _f, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "srb_grad_batch_create": (_i, [_vp, _vp, _i, _i, C.POINTER(_vp)]),
    "srb_grad_batch_units": (_i, [_vp]),
    "srb_grad_batch_run": (_i, [_vp, _vp, _i, _vp]),
    "srb_grad_batch_destroy": (None, [_vp]),
    "srb_pack_batch_create": (_i, [_vp, _vp, _i, C.POINTER(_vp)]),
    "srb_pack_batch_run": (_i, [_vp, _vp, _vp]),
    "srb_pack_batch_destroy": (None, [_vp]),
    "srb_set_programmatic_launch": (_i, [_vp, _i]),
    "srb_u8nchw_to_pair4": (_i, [_vp, _vp, _vp, _fp, _fp, _f, _i, _i, _i, _i, _vp]),
    "srb_pack_conv_pair4": (_i, [_vp, _vp, _vp, _i, _i, C.POINTER(_vp)]),
    "srb_resblock_create": (_i, [_vp, _vp, _vp, _vp, _vp, _i, C.POINTER(_vp)]),
    "srb_resblock_destroy": (None, [_vp]),
    "srb_resblock_supported": (_i, [_i, _i]),
    "srb_resblock_fwd": (_i, [_vp, _vp, _vp, _vp, _f, _i, _i, _i, _vp]),
    "srb_rowconv_create": (_i, [_vp, _vp, _vp, _i, C.POINTER(_vp)]),
    "srb_rowconv_destroy": (None, [_vp]),
    "srb_rowconv_fwd": (_i, [_vp, _vp, _vp, _vp, _vp, _f, _i, _i, _i, _i, _vp]),
    "srb_tail9_create": (_i, [_vp, _vp, _vp, _i, C.POINTER(_vp)]),
    "srb_tail9_destroy": (None, [_vp]),
    "srb_tail9_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "srb_optim_workspace_floats": (_i, []),
    "srb_grad_stats": (_i, [_vp, _vp, _sz, _vp, _vp, _vp, _vp]),
    "srb_adam_ema_step": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp, _vp, _vp, _f, _f, _f, _f, _f, _f, _f, _vp]),
    "srb_l1_partials_len": (_i, []),
    "srb_l1_loss_fwd": (_i, [_vp, _vp, _i, _vp, _sz, _f, _vp, _vp, _vp]),
    "srb_l1_loss_fwd_noalloc": (_i, [_vp, _vp, _i, _vp, _sz, _f, _vp, _vp]),
    "srb_l1_loss_bwd": (_i, [_vp, _vp, _i, _vp, _sz, _f, _f, _vp, _vp]),
    "srb_l1_loss_bwd_dev": (_i, [_vp, _vp, _i, _vp, _sz, _f, _f, _vp, _vp, _vp]),
    "srb_bn_stats": (_i, [_vp, _vp, _i, _vp]),
    "srb_bn_apply": (_i, [_vp, _vp, _i, _vp]),
    "srb_bn_bwd_reduce": (_i, [_vp, _vp, _i, _vp]),
    "srb_bn_bwd_apply": (_i, [_vp, _vp, _i, _vp]),
    "srb_mse_loss_fwd": (_i, [_vp, _vp, _i, _vp, _sz, _f, _vp, _vp, _vp]),
    "srb_mse_loss_bwd_dev": (_i, [_vp, _vp, _i, _vp, _sz, _f, _f, _vp, _vp, _vp]),
}

# probe hooks of the -DSRB_DEBUG build (include/srb200_debug.h): present only when SRB200_LIB points at libsrb200_dbg.so (tools/)
DEBUG_SIGNATURES = {
    "srb_debug_set": (_i, [_vp, _i, _i]),
    "srb_debug_trace": (_i, [_vp, _vp]),
    "srb_ubench_mma": (_i, [_vp, _i, _i, _i, _i, _i, _i, C.POINTER(C.c_longlong)]),
    "srb_ubench_mix": (_i, [_vp, _i, C.POINTER(_i), C.POINTER(_i), C.POINTER(_i), _i, _i, C.POINTER(C.c_longlong)]),
}
OPT_MAX_GRID, OPT_PATH = 1, 2
PATH_UNFUSED_TRUNK, PATH_NO_PDL, PATH_TILED_CONV1, PATH_TILED_TAIL, PATH_C8_HEAD = 8, 1024, 8192, 16384, 32768
PATH_LARGE_TILING, PATH_UNSORTED_UNITS, PATH_WS_LARGE_TILING, PATH_NO_PAIR = 65536, 131072, 262144, 1048576
PATH_GROUP_BOXES, PATH_PAIR4_HEAD, PATH_SPLIT_C1TAIL = 2097152, 4194304, 8388608

_lib = None
_lock = threading.Lock()


def lib():
    """Load libsrb200.so (once).  Raises if it has not been built — never falls back."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise ImportError(
                        f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                        "(nvcc, sm_100a). There is no CPU fallback.")
                l = C.CDLL(LIB_PATH)
                for name, (res, args) in SIGNATURES.items():
                    fn = getattr(l, name)  # AttributeError if the .so lacks a declared symbol
                    fn.restype = res
                    fn.argtypes = args
                for name, (res, args) in DEBUG_SIGNATURES.items():
                    fn = getattr(l, name, None)
                    if fn is not None:
                        fn.restype = res
                        fn.argtypes = args
                _lib = l
    return _lib


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = lib().srb_last_error().decode(errors="replace")
        raise RuntimeError(f"libsrb200 {what} failed ({rc}): {msg}")


class Context:
    """One srb_ctx per CUDA device (per process); owns nothing but staging buffers."""
    _by_device: dict[int, "Context"] = {}

    def __init__(self, device: int):
        self.device = int(device)
        h = _vp()
        check(lib().srb_create(self.device, C.byref(h)), "srb_create")
        self.handle = h
        self.num_sms = lib().srb_num_sms(h)

    @classmethod
    def get(cls, device: int) -> "Context":
        device = int(device)
        if device not in cls._by_device:
            cls._by_device[device] = Context(device)
        return cls._by_device[device]

    def check_device(self):
        """Synchronise and raise if a kernel reported a pipeline timeout."""
        check(lib().srb_check_device(self.handle), "srb_check_device")

    def poll(self):
        """Raise if a kernel of this device reported a pipeline timeout since the last check (no synchronisation: reads the
        host-mapped abort flag).  Called wherever results are handed to the caller."""
        check(lib().srb_poll_abort(self.handle), "srb_poll_abort")

    def set_option(self, key: int, value: int):
        """Host-side options (include/srb200.h: SRB_OPT_MAX_GRID, SRB_OPT_PATH)."""
        check(lib().srb_set_option(self.handle, key, value), "srb_set_option")

    def debug_set(self, key: int, value: int):
        """Probe hooks: only with the -DSRB_DEBUG build (SRB200_LIB=.../libsrb200_dbg.so); the product library has none."""
        fn = getattr(lib(), "srb_debug_set", None)
        if fn is None:
            raise RuntimeError("srb_debug_set exists only in libsrb200_dbg.so (build.build(debug=True), SRB200_LIB=...)")
        check(fn(self.handle, key, value), "srb_debug_set")


class ConvArgs(C.Structure):
    """srb_conv_args of include/srb200.h."""
    _fields_ = [("x", _vp), ("x_c8_total", _i), ("x_c8_off", _i),
                ("y", _vp), ("y_c8_total", _i), ("y_c8_off", _i),
                ("residual", _vp), ("res_c8_total", _i), ("res_c8_off", _i), ("res_nplanes", _i), ("res_scale", _f),
                ("res2", _vp), ("res2_c8_total", _i), ("res2_c8_off", _i), ("res2_scale", _f),
                ("out_scale", _f), ("lrelu_slope", _f), ("flags", _i),
                ("mask", _vp), ("mask_c8_total", _i), ("mask_c8_off", _i), ("mask_p0", _i), ("mask_p1", _i), ("mask_slope", _f),
                ("N", _i), ("H", _i), ("W", _i)]


class BNArgs(C.Structure):
    """srb_bn_args of include/srb200.h."""
    _fields_ = [("y", _vp), ("y_c8_total", _i), ("y_c8_off", _i),
                ("out", _vp), ("out_c8_total", _i), ("out_c8_off", _i),
                ("planes", _i), ("gamma", _vp), ("beta", _vp), ("sums", _vp),
                ("running_mean", _vp), ("running_var", _vp), ("num_batches_tracked", _vp),
                ("momentum", _f), ("eps", _f), ("slope", _f), ("act", _i), ("out_scale", _f),
                ("res", _vp), ("res_c8_total", _i), ("res_c8_off", _i), ("res_scale", _f),
                ("res2", _vp), ("res2_c8_total", _i), ("res2_c8_off", _i), ("res2_scale", _f),
                ("da", _vp), ("da_c8_total", _i), ("da_c8_off", _i), ("up_scale", _f),
                ("bsums", _vp), ("dgamma", _vp), ("dbeta", _vp), ("grad_scale", _f),
                ("N", _i), ("H", _i), ("W", _i)]


class GradJob(C.Structure):
    """srb_grad_job of include/srb200.h."""
    _fields_ = [("x", _vp), ("x_c8_total", _i), ("x_c8_off", _i), ("cin", _i),
                ("dy", _vp), ("dy_c8_total", _i), ("dy_c8_off", _i), ("cout", _i),
                ("dw", _vp), ("db", _vp), ("scale", _f), ("N", _i), ("H", _i), ("W", _i)]


class PackJob(C.Structure):
    """srb_pack_job of include/srb200.h."""
    _fields_ = [("w", _vp), ("w_src", _vp), ("src_o", _i), ("src_i", _i), ("o_off", _i), ("c_off", _i), ("transpose", _i),
                ("bias_src", _vp), ("c_lo", _i), ("c_n", _i), ("scale", _f)]


def f32_ptr(arr):
    """numpy float32 contiguous array -> void* (keeps no reference: caller holds arr)."""
    return arr.ctypes.data_as(_vp)
