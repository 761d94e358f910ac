"""CPU: the C-ABI library loads and exports every symbol include/rtb.h declares; without a GPU the
product path fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from raytracing_b200 import _cabi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "rtb.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(rtb_[a-z0-9_]+)\s*\(", txt)))


def test_exports_every_declared_symbol():
    lib = _cabi.load()
    syms = header_symbols()
    assert set(syms) == set(_cabi.EXPORTS)
    for s in syms:
        assert getattr(lib, s) is not None
    assert lib.rtb_abi_version() == _cabi.RTB_ABI_VERSION == 2


def test_struct_layout_matches_header():
    assert C.sizeof(_cabi.RtbSpline) == 3 * 8 + 4 * 4
    assert C.sizeof(_cabi.RtbParams) == 4 * 4 + 8 + 6 * 8 + 2 * 8
    assert C.sizeof(_cabi.RtbStats) == 7 * 8
    assert C.sizeof(_cabi.RtbSweepSpec) == 2 * 8 + 8 + 8 + 2 * 8   # int32 metric padded to 8


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_has_gpu(), reason="only meaningful without a GPU")
def test_no_cpu_fallback():
    lib = _cabi.load()
    ctx = C.c_void_p()
    assert lib.rtb_create(0, C.byref(ctx)) == _cabi.RTB_ERR_CUDA and not ctx
    import raytracing_b200 as rt
    with pytest.raises(_cabi.RtbError):
        rt.Engine(0)
    z, grd = rt.build_field("2")
    with pytest.raises(_cabi.RtbError):
        rt.trazar(6, z, grd, False, rt.DELTA_S, 91, "2")


def test_product_does_not_import_oracle():
    """The product package must never route through oracle/ (test infrastructure only)."""
    pkg = os.path.join(ROOT, "raytracing_b200")
    for dirpath, _d, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src, f


def test_library_was_built_from_this_tree():
    """rtb_build_id() is the digest of the CUDA sources the .so was compiled from (csrc/Makefile); it must equal the
    digest of the sources in the tree, so a stale prebuilt library cannot pass for the current code."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import build_id
    assert _cabi.load().rtb_build_id().decode() == build_id.digest()
    assert _cabi.load().rtb_kernel_id().decode() == build_id.kernel_digest()
