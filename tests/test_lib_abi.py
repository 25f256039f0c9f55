"""CPU tests of the drop-in boundary: the C-ABI library builds, loads without a GPU and exports every
symbol include/srb200.h declares; the ctypes binding covers all of them; errors are reported, not crashed."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built_lib():
    import __graft_entry__ as g
    g.build()
    from image_super_resolution_b200 import _lib
    return _lib


def header_symbols():
    src = open(os.path.join(ROOT, "include", "srb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(srb_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_expected_entry_points():
    syms = header_symbols()
    for must in ("srb_create", "srb_pack_conv", "srb_conv2d_fwd", "srb_model_pt_forward", "srb_model_pt_forward_host",
                 "srb_l1_loss_fwd", "srb_l1_loss_bwd", "srb_u8nchw_to_c8", "srb_last_error"):
        assert must in syms


def test_library_exports_every_declared_symbol(built_lib):
    l = ctypes.CDLL(built_lib.LIB_PATH)
    for s in header_symbols():
        assert hasattr(l, s), f"libsrb200.so lacks {s}"


def test_binding_covers_header(built_lib):
    assert sorted(built_lib.SIGNATURES) == header_symbols()
    built_lib.lib()  # resolves all


def test_product_library_carries_no_probe_hooks(built_lib):
    """The bottleneck probes / clock traces / micro-benchmarks live in the -DSRB_DEBUG build only (include/srb200_debug.h)."""
    l = ctypes.CDLL(built_lib.LIB_PATH)
    for s in built_lib.DEBUG_SIGNATURES:
        assert not hasattr(l, s), f"libsrb200.so exports the probe hook {s}"


def test_errors_are_reported_not_crashed(built_lib):
    import torch
    l = built_lib.lib()
    assert l.srb_version() >= 100
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu tests")
    h = ctypes.c_void_p()
    rc = l.srb_create(0, ctypes.byref(h))
    assert rc != 0 and h.value is None
    assert len(l.srb_last_error()) > 0
    assert l.srb_model_pt_workspace_bytes(0, 4, 4) == 0
    assert l.srb_model_pt_workspace_bytes(1, 8, 8) >= 64 * (16 + 3 * 128 + 64)


def test_product_has_no_oracle_or_cpu_fallback():
    """The product package must not import the oracle, and must refuse CPU devices."""
    pkg = os.path.join(ROOT, "image_super_resolution_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert "import oracle" not in src and "from oracle" not in src, fn
    import image_super_resolution_b200 as srb
    m = srb.load(os.path.join(ROOT, "tests", "golden", "model_pt_weights.npz"))
    with pytest.raises(RuntimeError):
        m.to("cpu")


def test_python_constants_mirror_the_header():
    """The option / path / layout constants of the ctypes binding are the header's #defines (a drifted bit would silently select
    another kernel family)."""
    from image_super_resolution_b200 import _lib
    src = open(os.path.join(ROOT, "include", "srb200.h")).read()
    defs = {m.group(1): int(m.group(2)) for m in re.finditer(r"#define\s+(SRB_[A-Z0-9_]+)\s+(\d+)\b", src)}
    for name, val in vars(_lib).items():
        if name.startswith(("PATH_", "OPT_")) and isinstance(val, int):
            assert defs.get("SRB_" + name) == val, (name, val, defs.get("SRB_" + name))
    assert defs["SRB_LAYOUT_NCHW_RGB"] == 0 and defs["SRB_LAYOUT_NHWC_BGR"] == 1 and defs["SRB_LAYOUT_NHWC_RGB"] == 2
    from image_super_resolution_b200.modelpt import ModelPT
    assert ModelPT._IN_LAYOUTS == {"nchw": 0, "hwc_bgr": 1, "hwc_rgb": 2}
