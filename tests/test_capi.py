"""CPU tests of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/phasm_ovl.h declares; without a GPU the product path fails loudly (no CPU fallback)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    with open(os.path.join(ROOT, "include", "phasm_ovl.h")) as f:
        src = re.sub(r"/\*.*?\*/", "", f.read(), flags=re.S)
    return sorted(set(re.findall(r"\b(ovl_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_expected_surface():
    syms = declared_symbols()
    for must in ("ovl_create", "ovl_destroy", "ovl_add_sequence", "ovl_overlaps", "ovl_last_error",
                 "ovl_add_read", "ovl_overlaps_device", "ovl_fetch_rows", "ovl_comm_init"):
        assert must in syms


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as g
    lib_path = g.build_library()
    lib = ctypes.CDLL(lib_path)
    for s in declared_symbols():
        assert hasattr(lib, s), "libphasm_ovl.so does not export %s" % s


def test_ctypes_binding_covers_header():
    from phasm_b200 import _capi
    assert sorted(_capi.SIGNATURES) == declared_symbols()
    _capi.load_library()
    assert ctypes.sizeof(_capi.OvlRow) == 20


def _has_gpu():
    try:
        lib = ctypes.CDLL("libcudart.so.12")
    except OSError:
        return False
    n = ctypes.c_int(0)
    return lib.cudaGetDeviceCount(ctypes.byref(n)) == 0 and n.value > 0


@pytest.mark.skipif(_has_gpu(), reason="a GPU is present")
def test_no_cpu_fallback_without_gpu():
    from phasm_b200._capi import OverlapperError
    from phasm_b200.overlapper import ExactOverlapper
    with pytest.raises(OverlapperError) as ei:
        ExactOverlapper()
    assert "no CPU fallback" in str(ei.value)


def test_product_package_never_imports_oracle():
    pkg = os.path.join(ROOT, "phasm_b200")
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                with open(os.path.join(dp, fn)) as f:
                    txt = f.read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), fn
                assert "overlap_oracle" not in txt, fn
